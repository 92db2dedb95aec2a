// Socket-address helpers: conversion between the public ccoip_socket_address_t and the kernel's
// sockaddr types, printing, ordering and hashing (reference: ccoip/src/cpp/ccoip_utils.cpp,
// ccoip/internal/ccoip_inet_utils.hpp).
#pragma once
#include <netinet/in.h>
#include <sys/socket.h>

#include <cstring>
#include <functional>
#include <string>

#include "pccl.h"

namespace pccl::net {

// Returns the sockaddr length, 0 on failure.
socklen_t to_sockaddr(const ccoip_socket_address_t &a, sockaddr_storage &out);
bool from_sockaddr(const sockaddr *sa, ccoip_socket_address_t &out);
std::string to_string(const ccoip_socket_address_t &a);
bool parse_address(const std::string &text, ccoip_socket_address_t &out); // "1.2.3.4:80" or "[::1]:80"
bool is_loopback(const ccoip_socket_address_t &a);
ccoip_socket_address_t any_v4(uint16_t port);
// Wildcard address of the same family as `like` ("::" listens dual-stack, see listen_tcp).
ccoip_socket_address_t any_like(const ccoip_socket_address_t &like, uint16_t port);

// Value type usable as map key.
struct AddrKey {
    uint8_t bytes[19]{}; // 1 family + 16 ip + 2 port
    explicit AddrKey(const ccoip_socket_address_t &a) {
        bytes[0] = a.inet.protocol == inetIPv4 ? 4 : 6;
        if (a.inet.protocol == inetIPv4)
            std::memcpy(bytes + 1, a.inet.ipv4.data, 4);
        else
            std::memcpy(bytes + 1, a.inet.ipv6.data, 16);
        bytes[17] = static_cast<uint8_t>(a.port >> 8);
        bytes[18] = static_cast<uint8_t>(a.port);
    }
    AddrKey() = default;
    bool operator==(const AddrKey &o) const { return std::memcmp(bytes, o.bytes, sizeof bytes) == 0; }
    bool operator<(const AddrKey &o) const { return std::memcmp(bytes, o.bytes, sizeof bytes) < 0; }
    ccoip_socket_address_t to_address() const {
        ccoip_socket_address_t a{};
        if (bytes[0] == 4) {
            a.inet.protocol = inetIPv4;
            std::memcpy(a.inet.ipv4.data, bytes + 1, 4);
        } else {
            a.inet.protocol = inetIPv6;
            std::memcpy(a.inet.ipv6.data, bytes + 1, 16);
        }
        a.port = static_cast<uint16_t>(bytes[17] << 8 | bytes[18]);
        return a;
    }
};

} // namespace pccl::net

template <> struct std::hash<pccl::net::AddrKey> {
    size_t operator()(const pccl::net::AddrKey &k) const noexcept {
        size_t h = 1469598103934665603ull;
        for (uint8_t b : k.bytes) h = (h ^ b) * 1099511628211ull;
        return h;
    }
};
