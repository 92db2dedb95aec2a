#include "queued_socket.hpp"

#include <unistd.h>

#include "../util/log.hpp"
#include "sockutil.hpp"

namespace pccl::net {

QueuedSocket::~QueuedSocket() {
    interrupt();
    join();
    if (fd_ >= 0) ::close(fd_);
}

bool QueuedSocket::connect() {
    if (fd_ >= 0) return false;
    fd_ = connect_tcp(remote_);
    if (fd_ < 0) return false;
    set_send_timeout(fd_, 10);
    set_keepalive(fd_, 60);
    open_ = true;
    return true;
}

bool QueuedSocket::run() {
    if (fd_ < 0 || rx_thread_.joinable()) return false;
    rx_thread_ = std::thread([this] { rx_loop(); });
    return true;
}

bool QueuedSocket::interrupt() {
    if (fd_ >= 0) ::shutdown(fd_, SHUT_RDWR);
    {
        std::lock_guard<std::mutex> lock(mu_);
        open_ = false;
    }
    cv_.notify_all();
    return true;
}

void QueuedSocket::join() {
    if (rx_thread_.joinable()) rx_thread_.join();
}

bool QueuedSocket::send_frame(const std::vector<uint8_t> &frame) {
    std::lock_guard<std::mutex> lock(send_mu_);
    if (!open_) return false;
    return send_all(fd_, frame.data(), frame.size());
}

void QueuedSocket::rx_loop() {
    for (;;) {
        uint8_t head[10];
        if (!recv_all(fd_, head, 8)) break;
        uint64_t len = 0;
        for (int i = 0; i < 8; ++i) len = (len << 8) | head[i];
        if (len < 2 || len > wire::kMaxControlFrame) break;
        if (!recv_all(fd_, head + 8, 2)) break;
        Frame f;
        f.id = static_cast<uint16_t>(head[8] << 8 | head[9]);
        f.payload.resize(len - 2);
        if (!f.payload.empty() && !recv_all(fd_, f.payload.data(), f.payload.size())) break;
        if (observer_) observer_(f.id, f.payload);
        {
            std::lock_guard<std::mutex> lock(mu_);
            inbox_.push_back(std::move(f));
            arrivals_.fetch_add(1, std::memory_order_release);
        }
        cv_.notify_all();
    }
    {
        std::lock_guard<std::mutex> lock(mu_);
        open_ = false;
    }
    cv_.notify_all();
}

} // namespace pccl::net
