#include "blocking_socket.hpp"

#include <sys/socket.h>
#include <unistd.h>

#include <cerrno>

#include "../util/log.hpp"
#include "sockutil.hpp"

namespace pccl::net {

bool BlockingSocket::connect(int timeout_ms) {
    if (fd_ >= 0) return false;
    fd_ = connect_tcp(remote_, timeout_ms);
    return fd_ >= 0;
}

void BlockingSocket::close() {
    if (fd_ >= 0) {
        // give queued bytes a chance to leave before the descriptor dies
        linger lg{1, 5};
        setsockopt(fd_, SOL_SOCKET, SO_LINGER, &lg, sizeof lg);
        ::shutdown(fd_, SHUT_RDWR);
        ::close(fd_);
        fd_ = -1;
    }
}

bool BlockingSocket::is_open() const { return peer_alive(fd_); }

bool BlockingSocket::send_raw(const void *data, size_t n) { return fd_ >= 0 && send_all(fd_, data, n); }
bool BlockingSocket::recv_raw(void *data, size_t n) { return fd_ >= 0 && recv_all(fd_, data, n); }

bool BlockingSocket::recv_frame(uint16_t &id, std::vector<uint8_t> &payload, int timeout_ms) {
    if (fd_ < 0) return false;
    if (timeout_ms >= 0) {
        if (wait_readable(fd_, timeout_ms) <= 0) return false;
    }
    uint8_t head[10];
    if (!recv_all(fd_, head, 8)) return false;
    uint64_t len = 0;
    for (int i = 0; i < 8; ++i) len = (len << 8) | head[i];
    if (len < 2 || len > wire::kMaxControlFrame) return false;
    if (!recv_all(fd_, head + 8, 2)) return false;
    id = static_cast<uint16_t>(head[8] << 8 | head[9]);
    payload.resize(len - 2);
    return payload.empty() || recv_all(fd_, payload.data(), payload.size());
}

Listener::~Listener() {
    interrupt();
    join();
}

bool Listener::listen() {
    if (fd_ >= 0) return false;
    fd_ = listen_tcp(addr_, bump_, port_);
    if (fd_ < 0) PLOG(ERR) << "cannot bind " << to_string(addr_) << ": " << strerror(errno);
    return fd_ >= 0;
}

bool Listener::run_async() {
    if (fd_ < 0 || running_.exchange(true)) return false;
    thread_ = std::thread([this] {
        while (!stop_) {
            const int rc = wait_readable(fd_, 200);
            if (rc == 0) continue;
            if (rc < 0) break;
            sockaddr_storage ss;
            socklen_t len = sizeof ss;
            const int cfd = ::accept4(fd_, reinterpret_cast<sockaddr *>(&ss), &len, SOCK_CLOEXEC);
            if (cfd < 0) {
                if (errno == EINTR || errno == EAGAIN) continue;
                break;
            }
            tune_stream_socket(cfd);
            ccoip_socket_address_t peer{};
            from_sockaddr(reinterpret_cast<sockaddr *>(&ss), peer);
            if (cb_)
                cb_(std::make_unique<BlockingSocket>(cfd, peer));
            else
                ::close(cfd);
        }
    });
    return true;
}

bool Listener::interrupt() {
    stop_ = true;
    return true;
}

void Listener::join() {
    if (thread_.joinable()) thread_.join();
    if (fd_ >= 0) {
        ::close(fd_);
        fd_ = -1;
    }
}

} // namespace pccl::net
