// Same-box operation sequencer: gives the peers of one ring that live on one host a common, total order of their
// device-plane collectives without a round trip to the master.
//
// The reference orders nothing locally: every all-reduce is a vote at the master (C2M CollectiveCommsInitiate ->
// M2C CollectiveCommsCommence, ccoip_client_handler.cpp:1226-1246), and the master's sequence number doubles as the
// order.  On loopback TCP that is ~100 us per operation -- irrelevant for a 600 MB reduction over a WAN, dominant for a
// 1 KB one over NVLink (option 4).  Under a lease (see client.cpp) peers skip the vote; what they
// still need is agreement on the ORDER in which concurrently issued operations (different tags, different threads) run,
// because two collective kernels that each wait for their twin on the other GPUs must be started in the same order
// everywhere.  That agreement lives here: an append-only log in a shared-memory segment (memfd, passed over the
// rendezvous socket).  The first peer to ask for operation (tag, k) appends it; everybody else finds it at the same
// index.  Lock-free (a peer that dies can never hold a lock), wait-free for readers of published entries.
//
// Layout of the segment: Header, then `capacity` entries used as a ring (entry of index i lives in slot i % capacity).
// An index is only ever reused after `capacity` further operations; collectives are, well, collective -- no peer can
// be more than its in-flight limit behind -- so capacity 4096 is ample.
#pragma once
#include <atomic>
#include <chrono>
#include <cstdint>
#include <cstring>
#include <unordered_map>

namespace pccl::device {

class Sequencer {
public:
    static constexpr uint64_t kMagic = 0x5043434c53455131ull; // "PCCLSEQ1"
    static constexpr uint32_t kCapacity = 4096;

    struct Entry {
        std::atomic<uint64_t> published; // index + 1 once tag/k are valid
        uint64_t tag, k;
        uint64_t pad;
    };
    struct Header {
        uint64_t magic;
        uint64_t base;                 // first flag epoch of this log, see flag_epoch()
        std::atomic<uint64_t> tail;    // next free index
        uint64_t capacity;
        uint8_t pad[32];
    };
    static_assert(sizeof(Entry) == 32 && sizeof(Header) == 64, "shared layout");
    static constexpr size_t bytes() { return sizeof(Header) + size_t(kCapacity) * sizeof(Entry); }

    // Flag epochs must grow from one log to the next (flag values left in a control page by an earlier ring must never
    // satisfy a wait of a later one): the creator stamps the log with a clock all processes of this host share.
    // 8192 epochs per millisecond is far more than any peer can issue.
    static uint64_t fresh_base() {
        timespec ts{};
        clock_gettime(CLOCK_MONOTONIC, &ts);
        const uint64_t ms = static_cast<uint64_t>(ts.tv_sec) * 1000ull + static_cast<uint64_t>(ts.tv_nsec) / 1000000ull;
        return ms * 8192ull;
    }

    // creator side: `mem` is a zero-filled region of bytes() bytes
    static void format(void *mem, uint64_t base) {
        auto *h = static_cast<Header *>(mem);
        h->base = base;
        h->capacity = kCapacity;
        h->tail.store(0, std::memory_order_relaxed);
        std::atomic_thread_fence(std::memory_order_release);
        h->magic = kMagic;
    }

    Sequencer() = default;
    bool attach(void *mem) {
        auto *h = static_cast<Header *>(mem);
        if (!h || h->magic != kMagic || h->capacity != kCapacity) return false;
        header_ = h;
        entries_ = reinterpret_cast<Entry *>(h + 1);
        cursor_ = 0;
        pending_.clear();
        return true;
    }
    void detach() {
        header_ = nullptr;
        entries_ = nullptr;
        pending_.clear();
    }
    bool attached() const { return header_ != nullptr; }
    uint64_t base() const { return header_ ? header_->base : 0; }
    uint64_t flag_epoch(uint64_t index) const { return base() + index; }

    // The position of operation (tag, k) in the common order; appends it when nobody has yet.  Not thread-safe: callers
    // serialise (one mutex per process).  Returns false when a claimed entry never gets published (its writer died) --
    // the lease is then dead anyway.
    bool order_of(uint64_t tag, uint64_t k, uint64_t &index, int timeout_ms = 2000) {
        if (!header_) return false;
        const Key key{tag, k};
        for (;;) {
            if (const auto it = pending_.find(key); it != pending_.end()) {
                index = it->second;
                pending_.erase(it);
                return true;
            }
            const uint64_t tail = header_->tail.load(std::memory_order_acquire);
            if (cursor_ < tail) {
                if (!scan(tail, timeout_ms)) return false;
                continue;
            }
            // nothing new: try to append at `tail`; succeeds only if nobody appended since the scan above
            uint64_t expect = tail;
            if (header_->tail.compare_exchange_strong(expect, tail + 1, std::memory_order_acq_rel)) {
                Entry &e = entries_[tail % kCapacity];
                e.tag = tag;
                e.k = k;
                e.published.store(tail + 1, std::memory_order_release);
                cursor_ = tail + 1;
                index = tail;
                return true;
            }
        }
    }

private:
    struct Key {
        uint64_t tag, k;
        bool operator==(const Key &o) const { return tag == o.tag && k == o.k; }
    };
    struct KeyHash {
        size_t operator()(const Key &x) const { return std::hash<uint64_t>()(x.tag * 0x9E3779B97F4A7C15ull ^ x.k); }
    };

    bool scan(uint64_t tail, int timeout_ms) {
        for (; cursor_ < tail; ++cursor_) {
            Entry &e = entries_[cursor_ % kCapacity];
            const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
            while (e.published.load(std::memory_order_acquire) != cursor_ + 1) {
                if (std::chrono::steady_clock::now() > deadline) return false;
#if defined(__x86_64__)
                __builtin_ia32_pause();
#endif
            }
            pending_.emplace(Key{e.tag, e.k}, cursor_);
        }
        return true;
    }

    Header *header_ = nullptr;
    Entry *entries_ = nullptr;
    uint64_t cursor_ = 0;                               // entries below it were examined
    std::unordered_map<Key, uint64_t, KeyHash> pending_; // seen in the log, not yet asked for by this process
};

} // namespace pccl::device
