// Single-threaded epoll TCP server that reassembles CCoIP control frames
// ([u64 BE length][u16 id][payload]) per client and hands complete frames to a callback.
//
// Behavioural model: reference tinysockets ServerSocket (tinysockets/src/server_socket.cpp:99-488):
// one loop thread owns all state, callbacks run on that thread, interrupt() is idempotent, a
// kicked client still produces exactly one close callback (delivered after the current
// dispatch returns, never re-entrantly).  Built on epoll + eventfd instead of libuv because
// the target is Linux only.
#pragma once
#include <atomic>
#include <deque>
#include <functional>
#include <mutex>
#include <span>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../wire/packets.hpp"
#include "addr.hpp"

namespace pccl::net {

class EpollServer {
public:
    using ReadCallback = std::function<void(const ccoip_socket_address_t &, std::span<const uint8_t>)>;
    using AddrCallback = std::function<void(const ccoip_socket_address_t &)>;

    explicit EpollServer(const ccoip_socket_address_t &listen_address, bool bump_port = false);
    ~EpollServer();

    void on_read(ReadCallback cb) { read_cb_ = std::move(cb); }   // frame = [u16 id][payload]
    void on_join(AddrCallback cb) { join_cb_ = std::move(cb); }
    void on_close(AddrCallback cb) { close_cb_ = std::move(cb); }

    bool listen();
    bool run_async();
    bool interrupt();
    void join();

    uint16_t port() const { return bound_port_; }
    std::thread::id loop_thread_id() const { return loop_thread_.get_id(); }

    // Both may be called from the loop thread (inside callbacks) or from any other thread.
    bool send_raw(const ccoip_socket_address_t &client, std::vector<uint8_t> bytes);
    template <typename P> bool send_packet(const ccoip_socket_address_t &client, const P &pkt) {
        return send_raw(client, wire::encode_frame(pkt));
    }
    bool close_client(const ccoip_socket_address_t &client);

private:
    struct Client {
        int fd = -1;
        ccoip_socket_address_t addr{};
        std::vector<uint8_t> rx;         // partial frame bytes
        std::deque<std::vector<uint8_t>> tx; // pending writes
        size_t tx_off = 0;
        bool closing = false;
        bool want_write = false;
        bool corked = false; // has frames whose flush is deferred to the end of the running callback
    };

    void loop();
    void accept_ready();
    void client_readable(Client &c);
    void flush(Client &c);
    void drop(int fd);
    void update_interest(Client &c);

    ccoip_socket_address_t listen_addr_;
    bool bump_;
    uint16_t bound_port_ = 0;
    int listen_fd_ = -1, epoll_fd_ = -1, wake_fd_ = -1;
    std::thread loop_thread_;
    std::atomic<bool> running_{false}, stop_{false}, listening_{false};

    std::recursive_mutex mu_; // guards clients_/by_addr_/pending_close_
    std::unordered_map<int, Client> clients_;
    std::unordered_map<AddrKey, int> by_addr_;
    std::vector<int> pending_close_;
    // Frames sent from inside a callback are written when the callback returns: a consensus that answers every
    // client with two or three packets costs one writev per client instead of one send per packet, and the client
    // wakes up once.  Only the loop thread touches these two (under mu_).
    int cork_depth_ = 0;
    std::vector<int> corked_;
    void uncork();
    struct Cork;

    ReadCallback read_cb_;
    AddrCallback join_cb_, close_cb_;
};

} // namespace pccl::net
