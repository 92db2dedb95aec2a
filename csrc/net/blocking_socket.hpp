// Blocking client socket speaking CCoIP control frames, plus an accept-thread listener.
// Behaviour model: reference tinysockets BlockingIOSocket / BlockingIOServerSocket
// (tinysockets/src/blocking_client_socket.cpp:9-245, blocking_server_socket.cpp:65-202).
#pragma once
#include <atomic>
#include <functional>
#include <memory>
#include <optional>
#include <thread>
#include <vector>

#include "../wire/packets.hpp"
#include "addr.hpp"

namespace pccl::net {

class BlockingSocket {
public:
    explicit BlockingSocket(const ccoip_socket_address_t &remote) : remote_(remote) {}
    BlockingSocket(int adopted_fd, const ccoip_socket_address_t &remote) : remote_(remote), fd_(adopted_fd) {}
    ~BlockingSocket() { close(); }
    BlockingSocket(const BlockingSocket &) = delete;
    BlockingSocket &operator=(const BlockingSocket &) = delete;

    bool connect(int timeout_ms = 15000);
    void close();
    bool is_open() const;
    int fd() const { return fd_; }
    int release() { // hand the descriptor to another owner
        const int f = fd_;
        fd_ = -1;
        return f;
    }
    const ccoip_socket_address_t &remote() const { return remote_; }

    template <typename P> bool send_packet(const P &pkt) {
        const auto frame = wire::encode_frame(pkt);
        return send_raw(frame.data(), frame.size());
    }
    // timeout_ms < 0 waits forever; no_wait returns immediately when nothing is readable.
    template <typename P> std::optional<P> recv_packet(int timeout_ms = -1) {
        std::vector<uint8_t> payload;
        uint16_t id = 0;
        if (!recv_frame(id, payload, timeout_ms)) return std::nullopt;
        if (id != P::id) return std::nullopt;
        wire::Reader r(payload.data(), payload.size());
        P pkt;
        if (!pkt.read(r)) return std::nullopt;
        return pkt;
    }
    bool recv_frame(uint16_t &id, std::vector<uint8_t> &payload, int timeout_ms = -1);
    bool send_raw(const void *data, size_t n);
    bool recv_raw(void *data, size_t n);

private:
    ccoip_socket_address_t remote_;
    int fd_ = -1;
};

// Accept loop on its own thread; every accepted connection is handed to the callback as an
// adopted BlockingSocket.
class Listener {
public:
    using JoinCallback = std::function<void(std::unique_ptr<BlockingSocket>)>;
    Listener(const ccoip_socket_address_t &addr, bool bump_port) : addr_(addr), bump_(bump_port) {}
    ~Listener();
    void on_join(JoinCallback cb) { cb_ = std::move(cb); }
    bool listen();
    bool run_async();
    bool interrupt();
    void join();
    uint16_t port() const { return port_; }

private:
    ccoip_socket_address_t addr_;
    bool bump_;
    uint16_t port_ = 0;
    int fd_ = -1;
    std::atomic<bool> stop_{false}, running_{false};
    std::thread thread_;
    JoinCallback cb_;
};

} // namespace pccl::net
