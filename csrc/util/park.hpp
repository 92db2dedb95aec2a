// One-word wait/wake primitive on top of the Linux futex call.
//
// A Parker is a 32-bit state word that one side publishes into and another side sleeps on.  It is
// what a send ticket or a "queue became non-empty" signal needs, at the cost of one atomic and no
// mutex: the waker stores the new value and issues FUTEX_WAKE only when somebody declared itself
// asleep; the sleeper re-checks the word around FUTEX_WAIT, so a wake between the check and the
// sleep is never lost (the kernel compares the word before blocking).
//
// Reference counterpart: third_party/threadpark (tparkBeginPark / tparkWait / tparkWake), which the
// reference's multiplexed socket uses for its TX thread and per-send done handles.
#pragma once
#include <atomic>
#include <cerrno>
#include <chrono>
#include <climits>
#include <cstdint>

#include <linux/futex.h>
#include <sys/syscall.h>
#include <time.h>
#include <unistd.h>

namespace pccl {

class Parker {
public:
    explicit Parker(uint32_t initial = 0) : word_(initial) {}

    uint32_t load() const { return word_.load(std::memory_order_acquire); }

    // Publishes `value` and wakes every sleeper.
    void publish(uint32_t value) {
        // store -> load of two different words: sequential consistency is what keeps the sleeper's
        // "announce, then check" and this "publish, then look for sleepers" from both missing each other
        word_.store(value, std::memory_order_seq_cst);
        if (sleepers_.load(std::memory_order_seq_cst) != 0) futex(FUTEX_WAKE_PRIVATE, INT_MAX, nullptr);
    }

    // Changes the word (by one) and wakes every sleeper: "something happened, look again".
    void bump() {
        word_.fetch_add(1, std::memory_order_seq_cst);
        if (sleepers_.load(std::memory_order_seq_cst) != 0) futex(FUTEX_WAKE_PRIVATE, INT_MAX, nullptr);
    }

    // Sleeps while the word equals `while_value`, at most `timeout_ms` (< 0: no limit).  Returns the word.
    uint32_t wait_while(uint32_t while_value, int timeout_ms) {
        uint32_t cur = load();
        if (cur != while_value) return cur;
        const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms < 0 ? 0 : timeout_ms);
        sleepers_.fetch_add(1, std::memory_order_seq_cst);
        while ((cur = word_.load(std::memory_order_seq_cst)) == while_value) {
            timespec ts{};
            timespec *tsp = nullptr;
            if (timeout_ms >= 0) {
                const auto left = deadline - std::chrono::steady_clock::now();
                if (left <= std::chrono::nanoseconds(0)) break;
                const auto ns = std::chrono::duration_cast<std::chrono::nanoseconds>(left).count();
                ts.tv_sec = static_cast<time_t>(ns / 1000000000LL);
                ts.tv_nsec = static_cast<long>(ns % 1000000000LL);
                tsp = &ts;
            }
            // returns at once with EAGAIN when the word already differs
            futex(FUTEX_WAIT_PRIVATE, static_cast<int>(while_value), tsp);
        }
        sleepers_.fetch_sub(1, std::memory_order_acq_rel);
        return cur;
    }

private:
    long futex(int op, int value, const timespec *timeout) {
        return syscall(SYS_futex, reinterpret_cast<uint32_t *>(&word_), op, value, timeout, nullptr, 0);
    }
    static_assert(sizeof(std::atomic<uint32_t>) == sizeof(uint32_t), "futex needs a plain 32-bit word");
    std::atomic<uint32_t> word_;
    std::atomic<uint32_t> sleepers_{0};
};

} // namespace pccl
