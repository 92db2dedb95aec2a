#include "epoll_server.hpp"

#include <sys/uio.h>

#include <sys/epoll.h>
#include <sys/eventfd.h>
#include <unistd.h>

#include <cerrno>
#include <csignal>

#include "../util/log.hpp"
#include "sockutil.hpp"

namespace pccl::net {

// RAII for "a callback is running on the loop thread"
struct EpollServer::Cork {
    EpollServer &s;
    explicit Cork(EpollServer &server) : s(server) {
        std::lock_guard<std::recursive_mutex> lock(s.mu_);
        ++s.cork_depth_;
    }
    ~Cork() {
        std::lock_guard<std::recursive_mutex> lock(s.mu_);
        if (--s.cork_depth_ == 0) s.uncork();
    }
};


EpollServer::EpollServer(const ccoip_socket_address_t &listen_address, bool bump_port)
    : listen_addr_(listen_address), bump_(bump_port) {
    ::signal(SIGPIPE, SIG_IGN);
}

EpollServer::~EpollServer() {
    interrupt();
    join();
    if (listen_fd_ >= 0) ::close(listen_fd_);
    if (epoll_fd_ >= 0) ::close(epoll_fd_);
    if (wake_fd_ >= 0) ::close(wake_fd_);
}

bool EpollServer::listen() {
    if (listening_.exchange(true)) return false;
    listen_fd_ = listen_tcp(listen_addr_, bump_, bound_port_);
    if (listen_fd_ < 0) {
        PLOG(ERR) << "cannot bind " << to_string(listen_addr_) << ": " << strerror(errno);
        listening_ = false;
        return false;
    }
    set_nonblocking(listen_fd_, true);
    epoll_fd_ = epoll_create1(EPOLL_CLOEXEC);
    wake_fd_ = eventfd(0, EFD_NONBLOCK | EFD_CLOEXEC);
    epoll_event ev{};
    ev.events = EPOLLIN;
    ev.data.fd = listen_fd_;
    epoll_ctl(epoll_fd_, EPOLL_CTL_ADD, listen_fd_, &ev);
    ev.data.fd = wake_fd_;
    epoll_ctl(epoll_fd_, EPOLL_CTL_ADD, wake_fd_, &ev);
    return true;
}

bool EpollServer::run_async() {
    // a stopped server stays stopped (reference: tinysockets/tests/test_server_socket.cpp test_run_async_after_interrupt): its
    // listening socket is gone, and re-assigning a thread that was never joined would terminate the process
    if (!listening_ || stop_ || running_.exchange(true)) return false;
    loop_thread_ = std::thread([this] { loop(); });
    return true;
}

bool EpollServer::interrupt() {
    if (!running_) return listening_.load() || true;
    stop_ = true;
    const uint64_t one = 1;
    if (wake_fd_ >= 0) (void) !::write(wake_fd_, &one, sizeof one);
    return true;
}

void EpollServer::join() {
    if (loop_thread_.joinable()) loop_thread_.join();
}

void EpollServer::update_interest(Client &c) {
    epoll_event ev{};
    ev.events = EPOLLIN | EPOLLRDHUP | (c.want_write ? static_cast<uint32_t>(EPOLLOUT) : 0u);
    ev.data.fd = c.fd;
    epoll_ctl(epoll_fd_, EPOLL_CTL_MOD, c.fd, &ev);
}

void EpollServer::accept_ready() {
    for (;;) {
        sockaddr_storage ss;
        socklen_t len = sizeof ss;
        const int fd = ::accept4(listen_fd_, reinterpret_cast<sockaddr *>(&ss), &len, SOCK_NONBLOCK | SOCK_CLOEXEC);
        if (fd < 0) {
            if (errno == EINTR) continue;
            return;
        }
        tune_stream_socket(fd);
        set_keepalive(fd, 60);
        Client c;
        c.fd = fd;
        from_sockaddr(reinterpret_cast<sockaddr *>(&ss), c.addr);
        const ccoip_socket_address_t addr = c.addr;
        {
            std::lock_guard<std::recursive_mutex> lock(mu_);
            by_addr_[AddrKey(addr)] = fd;
            clients_.emplace(fd, std::move(c));
        }
        epoll_event ev{};
        ev.events = EPOLLIN | EPOLLRDHUP;
        ev.data.fd = fd;
        epoll_ctl(epoll_fd_, EPOLL_CTL_ADD, fd, &ev);
        if (join_cb_) join_cb_(addr);
    }
}

void EpollServer::client_readable(Client &c) {
    Cork cork(*this);
    uint8_t buf[64 * 1024];
    for (;;) {
        const ssize_t k = ::recv(c.fd, buf, sizeof buf, 0);
        if (k > 0) {
            c.rx.insert(c.rx.end(), buf, buf + k);
            if (static_cast<size_t>(k) < sizeof buf) break;
            continue;
        }
        if (k == 0) {
            c.closing = true;
            break;
        }
        if (errno == EINTR) continue;
        if (errno == EAGAIN || errno == EWOULDBLOCK) break;
        c.closing = true;
        break;
    }
    // deliver every complete frame
    size_t off = 0;
    const int fd = c.fd;
    while (c.rx.size() - off >= 8) {
        uint64_t len = 0;
        for (int i = 0; i < 8; ++i) len = (len << 8) | c.rx[off + i];
        if (len < 2 || len > wire::kMaxControlFrame) {
            PLOG(WARN) << "client " << to_string(c.addr) << " sent an illegal frame length " << len;
            c.closing = true;
            break;
        }
        if (c.rx.size() - off - 8 < len) break;
        // copy out: the callback may kick clients / mutate the client table
        std::vector<uint8_t> frame(c.rx.begin() + off + 8, c.rx.begin() + off + 8 + len);
        off += 8 + len;
        const ccoip_socket_address_t addr = c.addr;
        if (read_cb_) read_cb_(addr, std::span<const uint8_t>(frame.data(), frame.size()));
        // the callback may have closed this client
        std::lock_guard<std::recursive_mutex> lock(mu_);
        auto it = clients_.find(fd);
        if (it == clients_.end() || it->second.closing) return;
    }
    if (off > 0) c.rx.erase(c.rx.begin(), c.rx.begin() + off);
}

void EpollServer::flush(Client &c) {
    while (!c.tx.empty()) {
        iovec iov[16];
        int n_iov = 0;
        for (auto it = c.tx.begin(); it != c.tx.end() && n_iov < 16; ++it, ++n_iov) {
            const size_t skip = n_iov == 0 ? c.tx_off : 0;
            iov[n_iov].iov_base = it->data() + skip;
            iov[n_iov].iov_len = it->size() - skip;
        }
        msghdr msg{};
        msg.msg_iov = iov;
        msg.msg_iovlen = static_cast<size_t>(n_iov);
        const ssize_t k = ::sendmsg(c.fd, &msg, MSG_NOSIGNAL);
        if (k < 0) {
            if (errno == EINTR) continue;
            if (errno == EAGAIN || errno == EWOULDBLOCK) break;
            c.closing = true;
            return;
        }
        size_t sent = static_cast<size_t>(k);
        while (sent > 0 && !c.tx.empty()) {
            const size_t left = c.tx.front().size() - c.tx_off;
            if (sent >= left) {
                sent -= left;
                c.tx.pop_front();
                c.tx_off = 0;
            } else {
                c.tx_off += sent;
                sent = 0;
            }
        }
    }
    const bool want = !c.tx.empty();
    if (want != c.want_write) {
        c.want_write = want;
        update_interest(c);
    }
}

// mu_ held, loop thread
void EpollServer::uncork() {
    std::vector<int> todo;
    todo.swap(corked_);
    for (int fd : todo) {
        auto it = clients_.find(fd);
        if (it == clients_.end()) continue;
        Client &c = it->second;
        c.corked = false;
        if (c.closing) continue;
        flush(c);
        if (c.closing) pending_close_.push_back(fd); // picked up by the loop's close pass of this iteration
    }
}

bool EpollServer::send_raw(const ccoip_socket_address_t &client, std::vector<uint8_t> bytes) {
    std::lock_guard<std::recursive_mutex> lock(mu_);
    const auto it = by_addr_.find(AddrKey(client));
    if (it == by_addr_.end()) return false;
    auto cit = clients_.find(it->second);
    if (cit == clients_.end() || cit->second.closing) return false;
    Client &c = cit->second;
    c.tx.push_back(std::move(bytes));
    if (cork_depth_ > 0 && std::this_thread::get_id() == loop_thread_.get_id()) {
        if (!c.corked) {
            c.corked = true;
            corked_.push_back(c.fd);
        }
        return true;
    }
    flush(c);
    if (c.closing) {
        pending_close_.push_back(c.fd);
        const uint64_t one = 1;
        (void) !::write(wake_fd_, &one, sizeof one);
        return false;
    }
    return true;
}

bool EpollServer::close_client(const ccoip_socket_address_t &client) {
    std::lock_guard<std::recursive_mutex> lock(mu_);
    const auto it = by_addr_.find(AddrKey(client));
    if (it == by_addr_.end()) return false;
    auto cit = clients_.find(it->second);
    if (cit == clients_.end()) return false;
    if (cit->second.closing) return true;
    cit->second.closing = true;
    ::shutdown(cit->second.fd, SHUT_RDWR);
    pending_close_.push_back(cit->second.fd);
    const uint64_t one = 1;
    (void) !::write(wake_fd_, &one, sizeof one);
    return true;
}

void EpollServer::drop(int fd) {
    ccoip_socket_address_t addr{};
    {
        std::lock_guard<std::recursive_mutex> lock(mu_);
        auto it = clients_.find(fd);
        if (it == clients_.end()) return;
        addr = it->second.addr;
        by_addr_.erase(AddrKey(addr));
        epoll_ctl(epoll_fd_, EPOLL_CTL_DEL, fd, nullptr);
        ::close(fd);
        clients_.erase(it);
    }
    Cork cork(*this);
    if (close_cb_) close_cb_(addr);
}

void EpollServer::loop() {
    std::vector<epoll_event> events(128);
    while (!stop_) {
        const int n = epoll_wait(epoll_fd_, events.data(), static_cast<int>(events.size()), 500);
        if (n < 0) {
            if (errno == EINTR) continue;
            break;
        }
        for (int i = 0; i < n && !stop_; ++i) {
            const int fd = events[i].data.fd;
            if (fd == listen_fd_) {
                accept_ready();
                continue;
            }
            if (fd == wake_fd_) {
                uint64_t v;
                (void) !::read(wake_fd_, &v, sizeof v);
                continue;
            }
            Client *c = nullptr;
            {
                std::lock_guard<std::recursive_mutex> lock(mu_);
                auto it = clients_.find(fd);
                if (it != clients_.end()) c = &it->second;
            }
            if (!c || c->closing) continue;
            if (events[i].events & EPOLLOUT) {
                std::lock_guard<std::recursive_mutex> lock(mu_);
                flush(*c);
            }
            if (events[i].events & (EPOLLIN | EPOLLRDHUP | EPOLLHUP | EPOLLERR)) client_readable(*c);
            {
                std::lock_guard<std::recursive_mutex> lock(mu_);
                auto it = clients_.find(fd);
                if (it != clients_.end() && it->second.closing) pending_close_.push_back(fd);
            }
        }
        // deliver deferred closes (kicks, EOFs, write failures); callbacks may queue more
        for (int rounds = 0; rounds < 64; ++rounds) {
            std::vector<int> todo;
            {
                std::lock_guard<std::recursive_mutex> lock(mu_);
                todo.swap(pending_close_);
            }
            if (todo.empty()) break;
            for (int fd : todo) drop(fd);
        }
    }
    // shutdown: close everything without callbacks into a half-dead owner
    std::lock_guard<std::recursive_mutex> lock(mu_);
    for (auto &[fd, c] : clients_) {
        epoll_ctl(epoll_fd_, EPOLL_CTL_DEL, fd, nullptr);
        shutdown_and_close(fd);
    }
    clients_.clear();
    by_addr_.clear();
    if (listen_fd_ >= 0) {
        ::close(listen_fd_);
        listen_fd_ = -1;
    }
    running_ = false;
}

} // namespace pccl::net
