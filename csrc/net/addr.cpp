#include "addr.hpp"

#include <arpa/inet.h>

namespace pccl::net {

socklen_t to_sockaddr(const ccoip_socket_address_t &a, sockaddr_storage &out) {
    std::memset(&out, 0, sizeof out);
    if (a.inet.protocol == inetIPv4) {
        auto *s = reinterpret_cast<sockaddr_in *>(&out);
        s->sin_family = AF_INET;
        s->sin_port = htons(a.port);
        std::memcpy(&s->sin_addr, a.inet.ipv4.data, 4);
        return sizeof(sockaddr_in);
    }
    if (a.inet.protocol == inetIPv6) {
        auto *s = reinterpret_cast<sockaddr_in6 *>(&out);
        s->sin6_family = AF_INET6;
        s->sin6_port = htons(a.port);
        std::memcpy(&s->sin6_addr, a.inet.ipv6.data, 16);
        return sizeof(sockaddr_in6);
    }
    return 0;
}

bool from_sockaddr(const sockaddr *sa, ccoip_socket_address_t &out) {
    out = ccoip_socket_address_t{};
    if (sa->sa_family == AF_INET) {
        const auto *s = reinterpret_cast<const sockaddr_in *>(sa);
        out.inet.protocol = inetIPv4;
        std::memcpy(out.inet.ipv4.data, &s->sin_addr, 4);
        out.port = ntohs(s->sin_port);
        return true;
    }
    if (sa->sa_family == AF_INET6) {
        const auto *s = reinterpret_cast<const sockaddr_in6 *>(sa);
        // v4-mapped addresses are reported as plain v4 so that identities are stable
        if (IN6_IS_ADDR_V4MAPPED(&s->sin6_addr)) {
            out.inet.protocol = inetIPv4;
            std::memcpy(out.inet.ipv4.data, s->sin6_addr.s6_addr + 12, 4);
        } else {
            out.inet.protocol = inetIPv6;
            std::memcpy(out.inet.ipv6.data, &s->sin6_addr, 16);
        }
        out.port = ntohs(s->sin6_port);
        return true;
    }
    return false;
}

std::string to_string(const ccoip_socket_address_t &a) {
    char buf[INET6_ADDRSTRLEN] = {0};
    if (a.inet.protocol == inetIPv4) {
        inet_ntop(AF_INET, a.inet.ipv4.data, buf, sizeof buf);
        return std::string(buf) + ":" + std::to_string(a.port);
    }
    inet_ntop(AF_INET6, a.inet.ipv6.data, buf, sizeof buf);
    return "[" + std::string(buf) + "]:" + std::to_string(a.port);
}

bool parse_address(const std::string &text, ccoip_socket_address_t &out) {
    out = ccoip_socket_address_t{};
    std::string host, port;
    if (!text.empty() && text[0] == '[') {
        const auto close = text.find(']');
        if (close == std::string::npos || close + 1 >= text.size() || text[close + 1] != ':') return false;
        host = text.substr(1, close - 1);
        port = text.substr(close + 2);
        out.inet.protocol = inetIPv6;
        if (inet_pton(AF_INET6, host.c_str(), out.inet.ipv6.data) != 1) return false;
    } else {
        const auto colon = text.rfind(':');
        if (colon == std::string::npos) return false;
        host = text.substr(0, colon);
        port = text.substr(colon + 1);
        out.inet.protocol = inetIPv4;
        if (inet_pton(AF_INET, host.c_str(), out.inet.ipv4.data) != 1) return false;
    }
    char *end = nullptr;
    const long p = strtol(port.c_str(), &end, 10);
    if (end == port.c_str() || *end != 0 || p < 0 || p > 65535) return false;
    out.port = static_cast<uint16_t>(p);
    return true;
}

bool is_loopback(const ccoip_socket_address_t &a) {
    if (a.inet.protocol == inetIPv4) return a.inet.ipv4.data[0] == 127;
    for (int i = 0; i < 15; ++i)
        if (a.inet.ipv6.data[i] != 0) return false;
    return a.inet.ipv6.data[15] == 1;
}

ccoip_socket_address_t any_v4(uint16_t port) {
    ccoip_socket_address_t a{};
    a.inet.protocol = inetIPv4;
    a.port = port;
    return a;
}

ccoip_socket_address_t any_like(const ccoip_socket_address_t &like, uint16_t port) {
    if (like.inet.protocol == inetIPv4) return any_v4(port);
    ccoip_socket_address_t a{};
    a.inet.protocol = inetIPv6; // all-zero address = "::"
    a.port = port;
    return a;
}

} // namespace pccl::net
