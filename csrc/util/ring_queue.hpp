// Bounded lock-free multi-producer / single-consumer queue (per-cell sequence numbers, after D. Vyukov's bounded
// MPMC queue restricted to one consumer).  The multiplexed socket's TX side uses it: reduce workers of any tag
// enqueue send requests, the socket's TX thread is the only consumer.
//
// Reference counterpart: tinysockets' vendored MPSC ring (mpsc/include/MPSCQueue.hpp, depth 1024) in front of its TX
// thread (multiplexed_socket.cpp:376-533).
#pragma once
#include <atomic>
#include <cstddef>
#include <cstdint>
#include <memory>
#include <new>
#include <utility>

namespace pccl {

template <typename T> class MpscRing {
public:
    // capacity is rounded up to a power of two
    explicit MpscRing(size_t capacity) {
        size_t n = 2;
        while (n < capacity) n <<= 1;
        mask_ = n - 1;
        cells_ = std::make_unique<Cell[]>(n);
        for (size_t i = 0; i < n; ++i) cells_[i].seq.store(i, std::memory_order_relaxed);
    }
    ~MpscRing() {
        T tmp;
        while (try_pop(tmp)) {
        }
    }
    MpscRing(const MpscRing &) = delete;
    MpscRing &operator=(const MpscRing &) = delete;

    // false when the ring is full; `value` is left untouched in that case
    bool try_push(T &&value) {
        size_t pos = head_.load(std::memory_order_relaxed);
        for (;;) {
            Cell &c = cells_[pos & mask_];
            const size_t seq = c.seq.load(std::memory_order_acquire);
            const intptr_t diff = static_cast<intptr_t>(seq) - static_cast<intptr_t>(pos);
            if (diff == 0) {
                if (head_.compare_exchange_weak(pos, pos + 1, std::memory_order_relaxed)) {
                    new (&c.storage) T(std::move(value));
                    c.seq.store(pos + 1, std::memory_order_release);
                    return true;
                }
            } else if (diff < 0) {
                return false; // the consumer has not freed this cell yet
            } else {
                pos = head_.load(std::memory_order_relaxed);
            }
        }
    }

    // single consumer only
    bool try_pop(T &out) {
        Cell &c = cells_[tail_ & mask_];
        const size_t seq = c.seq.load(std::memory_order_acquire);
        if (static_cast<intptr_t>(seq) - static_cast<intptr_t>(tail_ + 1) < 0) return false;
        T *slot = std::launder(reinterpret_cast<T *>(&c.storage));
        out = std::move(*slot);
        slot->~T();
        c.seq.store(tail_ + mask_ + 1, std::memory_order_release);
        ++tail_;
        return true;
    }

    size_t capacity() const { return mask_ + 1; }

private:
    struct Cell {
        std::atomic<size_t> seq;
        alignas(T) unsigned char storage[sizeof(T)];
    };
    std::unique_ptr<Cell[]> cells_;
    size_t mask_ = 0;
    alignas(64) std::atomic<size_t> head_{0}; // next position producers claim
    alignas(64) size_t tail_ = 0;             // next position the consumer reads
};

} // namespace pccl
