#include "sockutil.hpp"

#include <fcntl.h>
#include <netinet/tcp.h>
#include <poll.h>
#include <unistd.h>

#include <cerrno>

#include "../util/log.hpp"

namespace pccl::net {

int listen_tcp(const ccoip_socket_address_t &addr, bool bump, uint16_t &bound_port, int max_bumps) {
    ccoip_socket_address_t a = addr;
    for (int attempt = 0; attempt <= (bump ? max_bumps : 0); ++attempt) {
        const int family = a.inet.protocol == inetIPv4 ? AF_INET : AF_INET6;
        const int fd = ::socket(family, SOCK_STREAM | SOCK_CLOEXEC, 0);
        if (fd < 0) return -1;
        // No SO_REUSEADDR when bumping: two peers on one host must not share a port.
        int one = 1;
        if (!bump) setsockopt(fd, SOL_SOCKET, SO_REUSEADDR, &one, sizeof one);
        if (family == AF_INET6) { // an IPv6 wildcard listener also takes IPv4 peers (mapped addresses)
            int v6only = 0;
            setsockopt(fd, IPPROTO_IPV6, IPV6_V6ONLY, &v6only, sizeof v6only);
        }
        sockaddr_storage ss;
        const socklen_t len = to_sockaddr(a, ss);
        if (::bind(fd, reinterpret_cast<sockaddr *>(&ss), len) == 0 && ::listen(fd, 512) == 0) {
            sockaddr_storage actual;
            socklen_t alen = sizeof actual;
            getsockname(fd, reinterpret_cast<sockaddr *>(&actual), &alen);
            ccoip_socket_address_t got{};
            from_sockaddr(reinterpret_cast<sockaddr *>(&actual), got);
            bound_port = got.port;
            return fd;
        }
        ::close(fd);
        if (!bump || a.port == 0 || a.port == 65535) break;
        a.port++;
    }
    return -1;
}

void tune_stream_socket(int fd) {
    int one = 1;
    setsockopt(fd, IPPROTO_TCP, TCP_NODELAY, &one, sizeof one);
#ifdef TCP_QUICKACK
    setsockopt(fd, IPPROTO_TCP, TCP_QUICKACK, &one, sizeof one);
#endif
#ifdef SO_BUSY_POLL
    int busy_us = 50;
    setsockopt(fd, SOL_SOCKET, SO_BUSY_POLL, &busy_us, sizeof busy_us);
#endif
    int buf = 8 << 20;
    setsockopt(fd, SOL_SOCKET, SO_SNDBUF, &buf, sizeof buf);
    setsockopt(fd, SOL_SOCKET, SO_RCVBUF, &buf, sizeof buf);
}

void set_keepalive(int fd, int idle_seconds) {
    int one = 1;
    setsockopt(fd, SOL_SOCKET, SO_KEEPALIVE, &one, sizeof one);
    setsockopt(fd, IPPROTO_TCP, TCP_KEEPIDLE, &idle_seconds, sizeof idle_seconds);
    int intvl = 5, cnt = 3;
    setsockopt(fd, IPPROTO_TCP, TCP_KEEPINTVL, &intvl, sizeof intvl);
    setsockopt(fd, IPPROTO_TCP, TCP_KEEPCNT, &cnt, sizeof cnt);
}

void set_send_timeout(int fd, int seconds) {
    timeval tv{seconds, 0};
    setsockopt(fd, SOL_SOCKET, SO_SNDTIMEO, &tv, sizeof tv);
}

void set_nonblocking(int fd, bool on) {
    const int flags = fcntl(fd, F_GETFL, 0);
    fcntl(fd, F_SETFL, on ? (flags | O_NONBLOCK) : (flags & ~O_NONBLOCK));
}

int connect_tcp(const ccoip_socket_address_t &addr, int timeout_ms) {
    const int family = addr.inet.protocol == inetIPv4 ? AF_INET : AF_INET6;
    const int fd = ::socket(family, SOCK_STREAM | SOCK_CLOEXEC, 0);
    if (fd < 0) return -1;
    sockaddr_storage ss;
    const socklen_t len = to_sockaddr(addr, ss);
    set_nonblocking(fd, true);
    int rc = ::connect(fd, reinterpret_cast<sockaddr *>(&ss), len);
    if (rc != 0 && errno != EINPROGRESS) {
        ::close(fd);
        return -1;
    }
    if (rc != 0) {
        pollfd p{fd, POLLOUT, 0};
        do {
            rc = ::poll(&p, 1, timeout_ms);
        } while (rc < 0 && errno == EINTR);
        int err = 0;
        socklen_t elen = sizeof err;
        if (rc <= 0 || getsockopt(fd, SOL_SOCKET, SO_ERROR, &err, &elen) != 0 || err != 0) {
            ::close(fd);
            return -1;
        }
    }
    set_nonblocking(fd, false);
    tune_stream_socket(fd);
    return fd;
}

bool send_all(int fd, const void *data, size_t n) {
    const auto *p = static_cast<const uint8_t *>(data);
    while (n > 0) {
        const ssize_t k = ::send(fd, p, n, MSG_NOSIGNAL);
        if (k < 0) {
            if (errno == EINTR) continue;
            return false;
        }
        if (k == 0) return false;
        p += k;
        n -= static_cast<size_t>(k);
    }
    return true;
}

bool recv_all(int fd, void *data, size_t n) {
    auto *p = static_cast<uint8_t *>(data);
    while (n > 0) {
        const ssize_t k = ::recv(fd, p, n, 0);
        if (k < 0) {
            if (errno == EINTR) continue;
            return false;
        }
        if (k == 0) return false;
        p += k;
        n -= static_cast<size_t>(k);
    }
    return true;
}

int wait_readable(int fd, int timeout_ms) {
    pollfd p{fd, POLLIN, 0};
    int rc;
    do {
        rc = ::poll(&p, 1, timeout_ms);
    } while (rc < 0 && errno == EINTR);
    if (rc < 0) return -1;
    if (rc == 0) return 0;
    if (p.revents & POLLIN) return 1; // data (or EOF, which recv reports)
    return -1;
}

bool peer_alive(int fd) {
    if (fd < 0) return false;
    char c;
    const ssize_t k = ::recv(fd, &c, 1, MSG_PEEK | MSG_DONTWAIT);
    if (k > 0) return true;
    if (k == 0) return false;
    return errno == EAGAIN || errno == EWOULDBLOCK || errno == EINTR;
}

void shutdown_and_close(int fd) {
    if (fd < 0) return;
    ::shutdown(fd, SHUT_RDWR);
    ::close(fd);
}

} // namespace pccl::net
