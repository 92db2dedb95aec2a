// Thin helpers over BSD sockets shared by every transport class.
#pragma once
#include <cstddef>
#include <cstdint>

#include "addr.hpp"

namespace pccl::net {

// Creates a listening TCP socket on `addr`.  When `bump` is set a busy port is retried upwards
// (at most `max_bumps` times); the bound port is returned through `bound_port`.
// Reference behaviour: tinysockets/src/server_socket.cpp:113-163.
int listen_tcp(const ccoip_socket_address_t &addr, bool bump, uint16_t &bound_port, int max_bumps = 512);

// Blocking connect with the low-latency options the data plane wants.  Returns fd or -1.
int connect_tcp(const ccoip_socket_address_t &addr, int timeout_ms = 15000);

void tune_stream_socket(int fd);            // TCP_NODELAY, QUICKACK, busy-poll, big buffers
void set_keepalive(int fd, int idle_seconds);
void set_send_timeout(int fd, int seconds);
void set_nonblocking(int fd, bool on);

// Full-length blocking transfers; false on EOF / error.
bool send_all(int fd, const void *data, size_t n);
bool recv_all(int fd, void *data, size_t n);

// Waits until fd is readable.  >0 readable, 0 timeout, <0 error/hup.
int wait_readable(int fd, int timeout_ms);

// True while the peer has not closed (MSG_PEEK probe).
bool peer_alive(int fd);

void shutdown_and_close(int fd);

} // namespace pccl::net
