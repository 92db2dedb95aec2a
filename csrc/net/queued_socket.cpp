#include "queued_socket.hpp"

#include <sys/socket.h>
#include <unistd.h>

#include <algorithm>
#include <cerrno>
#include <cstring>

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
    // Buffered: the master answers a consensus with two or three small frames in one segment; one recv() takes
    // them all instead of three recv() calls (length, id, body) per frame.
    std::vector<uint8_t> buf(64 * 1024);
    size_t have = 0, pos = 0;
    auto need = [&](size_t n) { // at least n unread bytes in the buffer (n <= 10)
        while (have - pos < n) {
            if (pos > 0 && buf.size() - have < n) {
                std::memmove(buf.data(), buf.data() + pos, have - pos);
                have -= pos;
                pos = 0;
            }
            const ssize_t k = ::recv(fd_, buf.data() + have, buf.size() - have, 0);
            if (k < 0 && errno == EINTR) continue;
            if (k <= 0) return false;
            have += static_cast<size_t>(k);
        }
        return true;
    };
    for (;;) {
        if (!need(10)) break;
        const uint8_t *head = buf.data() + pos;
        uint64_t len = 0;
        for (int i = 0; i < 8; ++i) len = (len << 8) | head[i];
        if (len < 2 || len > wire::kMaxControlFrame) break;
        Frame f;
        f.id = static_cast<uint16_t>(head[8] << 8 | head[9]);
        pos += 10;
        f.payload.resize(len - 2);
        const size_t from_buffer = std::min(f.payload.size(), have - pos);
        if (from_buffer) std::memcpy(f.payload.data(), buf.data() + pos, from_buffer);
        pos += from_buffer;
        if (from_buffer < f.payload.size() && !recv_all(fd_, f.payload.data() + from_buffer, f.payload.size() - from_buffer)) break;
        if (pos == have) pos = have = 0;
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
