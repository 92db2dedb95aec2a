// Client side of the master link.  A receive thread parks complete frames in an ordered inbox;
// any number of consumer threads pick frames by packet type and, optionally, by a predicate on
// the decoded packet (for example "the Commence packet for MY tag") without disturbing the order
// of the frames they skip.  Behaviour model: reference tinysockets QueuedSocket
// (tinysockets/src/queued_client_socket.cpp:28-424).
#pragma once
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <cstdlib>
#include <functional>
#include <list>
#include <mutex>
#include <optional>
#include <thread>
#include <vector>

#include "../wire/packets.hpp"
#include "addr.hpp"

namespace pccl::net {

class QueuedSocket {
public:
    explicit QueuedSocket(const ccoip_socket_address_t &remote) : remote_(remote) {}
    ~QueuedSocket();
    QueuedSocket(const QueuedSocket &) = delete;
    QueuedSocket &operator=(const QueuedSocket &) = delete;

    bool connect();
    bool run(); // starts the receive thread
    // Called on the receive thread for every frame, in arrival order, before the frame is queued.
    using Observer = std::function<void(uint16_t id, const std::vector<uint8_t> &payload)>;
    void set_observer(Observer fn) { observer_ = std::move(fn); }
    bool interrupt();
    void join();
    bool is_open() const { return open_.load(); }

    template <typename P> bool send_packet(const P &pkt) {
        const auto frame = wire::encode_frame(pkt);
        return send_frame(frame);
    }

    template <typename P> std::optional<P> recv_packet(bool no_wait = false) {
        return recv_matching<P>([](const P &) { return true; }, no_wait);
    }

    // Returns the oldest queued packet of type P for which pred holds.  Blocks (unless no_wait)
    // until one arrives or the connection dies.
    template <typename P, typename Pred> std::optional<P> recv_matching(Pred pred, bool no_wait = false) {
        std::unique_lock<std::mutex> lock(mu_);
        // Answers from the master usually arrive within tens of microseconds: poll for a short
        // while before paying for a futex sleep and wake-up.
        auto spin_until = std::chrono::steady_clock::now() + std::chrono::microseconds(no_wait ? 0 : spin_budget_us());
        for (;;) {
            for (auto it = inbox_.begin(); it != inbox_.end(); ++it) {
                if (it->id != P::id) continue;
                wire::Reader r(it->payload.data(), it->payload.size());
                P pkt;
                if (!pkt.read(r)) {
                    inbox_.erase(it); // undecodable frame: drop it, report failure
                    return std::nullopt;
                }
                if (!pred(pkt)) continue;
                inbox_.erase(it);
                return pkt;
            }
            if (no_wait || !open_) return std::nullopt;
            if (std::chrono::steady_clock::now() < spin_until) {
                const size_t seen = arrivals_;
                lock.unlock();
                while (arrivals_.load(std::memory_order_acquire) == seen && std::chrono::steady_clock::now() < spin_until) {
#if defined(__x86_64__)
                    __builtin_ia32_pause();
#endif
                }
                lock.lock();
                continue;
            }
            cv_.wait(lock);
        }
    }

    // PCCL_SPIN_US overrides; by default only machines with cores to spare (16+) spin (on a small host
    // the spinning thread takes the core away from the thread it is waiting for).
    static int spin_budget_us() {
        static const int us = [] {
            if (const char *v = std::getenv("PCCL_SPIN_US")) return std::atoi(v);
            return std::thread::hardware_concurrency() >= 16 ? 60 : 0;
        }();
        return us;
    }

private:
    struct Frame {
        uint16_t id;
        std::vector<uint8_t> payload;
    };
    bool send_frame(const std::vector<uint8_t> &frame);
    void rx_loop();

    ccoip_socket_address_t remote_;
    int fd_ = -1;
    std::atomic<bool> open_{false};
    std::thread rx_thread_;
    std::mutex mu_, send_mu_;
    std::condition_variable cv_;
    std::list<Frame> inbox_;
    std::atomic<size_t> arrivals_{0}; // bumped for every queued frame; lets waiters poll without the lock
    Observer observer_;
};

} // namespace pccl::net
