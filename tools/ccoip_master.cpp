// Stand-alone master process (reference: ccoip_master/src/main.cpp:18-34): binds 0.0.0.0:48148 unless
// --listen-address / --port say otherwise, runs until SIGINT / SIGTERM.
#include <atomic>
#include <csignal>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>

#include "pccl.h"

static pcclMasterInstance_t *g_master = nullptr;

static void on_signal(int) {
    if (g_master) pcclInterruptMaster(g_master);
}

int main(int argc, char **argv) {
    ccoip_socket_address_t addr{};
    addr.inet.protocol = inetIPv4;
    addr.port = CCOIP_PROTOCOL_PORT_MASTER;
    for (int i = 1; i < argc; ++i) {
        const std::string arg = argv[i];
        if ((arg == "--port" || arg == "-p") && i + 1 < argc) {
            addr.port = static_cast<uint16_t>(atoi(argv[++i]));
        } else if (arg == "--listen-address" && i + 1 < argc) {
            unsigned a, b, c, d, port = addr.port;
            const int n = sscanf(argv[++i], "%u.%u.%u.%u:%u", &a, &b, &c, &d, &port);
            if (n < 4) {
                fprintf(stderr, "cannot parse address '%s'\n", argv[i]);
                return 2;
            }
            addr.inet.ipv4.data[0] = static_cast<uint8_t>(a);
            addr.inet.ipv4.data[1] = static_cast<uint8_t>(b);
            addr.inet.ipv4.data[2] = static_cast<uint8_t>(c);
            addr.inet.ipv4.data[3] = static_cast<uint8_t>(d);
            addr.port = static_cast<uint16_t>(port);
        } else if (arg == "--help" || arg == "-h") {
            printf("usage: ccoip_master [--listen-address a.b.c.d[:port]] [--port N]\n");
            return 0;
        }
    }
    pcclInit();
    if (pcclCreateMaster(addr, &g_master) != pcclSuccess) return 1;
    signal(SIGINT, on_signal);
    signal(SIGTERM, on_signal);
    if (pcclRunMaster(g_master) != pcclSuccess) {
        fprintf(stderr, "cannot start master on port %u\n", addr.port);
        return 1;
    }
    uint16_t port = 0;
    pcclExtMasterGetPort(g_master, &port);
    printf("ccoip_master listening on port %u\n", port);
    fflush(stdout);
    pcclMasterAwaitTermination(g_master);
    pcclDestroyMaster(g_master);
    return 0;
}
