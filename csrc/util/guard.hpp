// Small runtime assertions about who may be where.
//
//  * ThreadGuard: a piece of state that only one thread is allowed to touch (the master's whole state
//    machine lives on the epoll thread).  The first check() adopts the calling thread, later checks
//    from any other thread are a programming error and end the process with a message.
//    Reference counterpart: THREAD_GUARD, ccoip/internal_include/thread_guard.hpp:9-13.
//  * PhaseGuard: marks "this object is inside phase X" for the duration of a scope, so that other
//    entry points can refuse to run meanwhile (no topology change while reduces are in flight, ...).
//    Reference counterpart: guard_utils.hpp:9-21.
#pragma once
#include <atomic>
#include <thread>

#include "util/log.hpp"

namespace pccl {

class ThreadGuard {
public:
    void check(const char *what) {
        const std::thread::id me = std::this_thread::get_id();
        std::thread::id none{};
        if (owner_.compare_exchange_strong(none, me, std::memory_order_acq_rel)) return;
        if (none == me) return;
        PLOG(FATAL) << what << " was entered from a thread that does not own it";
    }
    // hand the state to whichever thread checks next (the owner thread has ended)
    void release() { owner_.store(std::thread::id{}, std::memory_order_release); }

private:
    std::atomic<std::thread::id> owner_{};
};

class PhaseGuard {
public:
    // `entered` is false when the phase was already active; the caller decides what that means
    explicit PhaseGuard(std::atomic<bool> &flag) : flag_(flag) {
        bool expected = false;
        entered_ = flag_.compare_exchange_strong(expected, true, std::memory_order_acq_rel);
    }
    ~PhaseGuard() {
        if (entered_) flag_.store(false, std::memory_order_release);
    }
    PhaseGuard(const PhaseGuard &) = delete;
    PhaseGuard &operator=(const PhaseGuard &) = delete;
    bool entered() const { return entered_; }

private:
    std::atomic<bool> &flag_;
    bool entered_;
};

} // namespace pccl
