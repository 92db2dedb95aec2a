// Several all-reduces in flight at once, three ways, with an in-process master when asked for.
//
//   ./concurrent_allreduce --master           (terminal 1: runs the master until Ctrl-C)
//   ./concurrent_allreduce 127.0.0.1:48148    (terminals 2..n: peers; they wait until two are there)
//
//  1. pcclAllReduceAsync on distinct tags + pcclAwaitAsyncReduce           (a tag is the identity of an operation
//     across peers: every peer must launch the same tags, in any order, and a tag can be in flight only once)
//  2. the same with uint8 min-max quantisation: a quarter of the bytes on the wire, bit-identical results everywhere
//  3. pcclAllReduceMultipleWithRetry: the library schedules a list of operations, at most `max_in_flight` at a time,
//     and starts the unfinished ones again on the shrunken ring when a peer is lost
// Counterpart of the reference's docs/md/06-Examples & Tutorials/05_ConcurrentAllReduce.md and
// tests/concurrent_reduce_test/main.cpp.
//
// Build:  g++ -std=c++17 -Iinclude examples/cpp/concurrent_allreduce.cpp -Lpccl_b200/lib -lpccl -Wl,-rpath,$PWD/pccl_b200/lib -pthread
#include <pccl.h>

#include <csignal>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#define CHECK(call)                                                                        \
    do {                                                                                   \
        const pcclResult_t rc_ = (call);                                                   \
        if (rc_ != pcclSuccess) {                                                          \
            std::fprintf(stderr, "%s failed with %d at line %d\n", #call, int(rc_), __LINE__); \
            std::exit(1);                                                                  \
        }                                                                                  \
    } while (0)

static pcclMasterInstance_t *g_master = nullptr;

static ccoip_socket_address_t parse(const std::string &text) {
    ccoip_socket_address_t a{};
    unsigned b[4] = {0, 0, 0, 0}, port = CCOIP_PROTOCOL_PORT_MASTER;
    if (std::sscanf(text.c_str(), "%u.%u.%u.%u:%u", &b[0], &b[1], &b[2], &b[3], &port) < 4) {
        std::fprintf(stderr, "expected a.b.c.d[:port], got '%s'\n", text.c_str());
        std::exit(2);
    }
    a.inet.protocol = inetIPv4;
    for (int i = 0; i < 4; ++i) a.inet.ipv4.data[i] = static_cast<uint8_t>(b[i]);
    a.port = static_cast<uint16_t>(port);
    return a;
}

static int run_master(const std::string &listen) {
    CHECK(pcclCreateMaster(parse(listen), &g_master));
    CHECK(pcclRunMaster(g_master));
    std::signal(SIGINT, [](int) { pcclInterruptMaster(g_master); });
    std::signal(SIGTERM, [](int) { pcclInterruptMaster(g_master); });
    uint16_t port = 0;
    CHECK(pcclExtMasterGetPort(g_master, &port));
    std::printf("master listening on port %u\n", port);
    std::fflush(stdout);
    CHECK(pcclMasterAwaitTermination(g_master));
    CHECK(pcclDestroyMaster(g_master));
    return 0;
}

int main(int argc, char **argv) {
    CHECK(pcclInit());
    if (argc > 1 && std::string(argv[1]) == "--master") return run_master(argc > 2 ? argv[2] : "0.0.0.0:48148");

    pcclCommCreateParams_t params{};
    params.master_address = parse(argc > 1 ? argv[1] : "127.0.0.1:48148");
    params.p2p_connection_pool_size = 4; // four sockets per neighbour: concurrent operations do not queue behind each other
    pcclComm_t *comm = nullptr;
    CHECK(pcclCreateCommunicator(&params, &comm));
    CHECK(pcclConnect(comm));
    int world = 0;
    for (;;) {
        CHECK(pcclGetAttribute(comm, PCCL_ATTRIBUTE_GLOBAL_WORLD_SIZE, &world));
        if (world >= 2) break;
        CHECK(pcclUpdateTopology(comm)); // admits whoever is waiting
    }

    constexpr int kOps = 8;
    constexpr size_t kCount = 1 << 18;
    std::vector<std::vector<float>> send(kOps, std::vector<float>(kCount)), recv(kOps, std::vector<float>(kCount));
    for (int t = 0; t < kOps; ++t)
        for (size_t i = 0; i < kCount; ++i) send[t][i] = float(t + 1) + float(i % 13) * 0.125f;

    // ---- 1. + 2. asynchronous operations, plain and quantised
    for (int quantised = 0; quantised < 2; ++quantised) {
        std::vector<pcclAsyncReduceOp_t> handles(kOps);
        for (int t = 0; t < kOps; ++t) {
            pcclReduceDescriptor_t d{};
            d.count = kCount, d.op = pcclSum, d.tag = uint64_t(t);
            d.src_descriptor.datatype = pcclFloat;
            if (quantised) d.quantization_options = {pcclUint8, pcclQuantMinMax};
            CHECK(pcclAllReduceAsync(send[t].data(), recv[t].data(), &d, comm, &handles[t]));
        }
        uint64_t tx = 0;
        for (int t = 0; t < kOps; ++t) {
            pcclReduceInfo_t info{};
            CHECK(pcclAwaitAsyncReduce(&handles[t], &info)); // pcclRankConnectionLost here = launch it again
            tx += info.tx_bytes;
            world = int(info.local_world_size);
        }
        double worst = 0;
        for (int t = 0; t < kOps; ++t)
            for (size_t i = 0; i < kCount; i += 101) {
                const double err = std::abs(double(recv[t][i]) - double(send[t][i]) * world);
                worst = err > worst ? err : worst;
            }
        std::printf("%s: %d operations of %zu floats between %d peers, %llu bytes sent, largest error %.4f\n", quantised ? "uint8 min-max" : "plain fp32",
                    kOps, kCount, world, static_cast<unsigned long long>(tx), worst);
    }

    // ---- 3. a list of operations, scheduled and retried by the library
    std::vector<pcclReduceOpDescriptor_t> list(kOps);
    for (int t = 0; t < kOps; ++t) {
        list[t].sendbuf = send[t].data(), list[t].recvbuf = recv[t].data();
        list[t].descriptor = {};
        list[t].descriptor.count = kCount, list[t].descriptor.op = pcclAvg, list[t].descriptor.tag = uint64_t(100 + t);
        list[t].descriptor.src_descriptor.datatype = pcclFloat;
    }
    pcclReduceInfo_t total{};
    const pcclResult_t rc = pcclAllReduceMultipleWithRetry(list.data(), list.size(), comm, &total, /*max_in_flight=*/4);
    if (rc == pcclTooFewPeers) {
        std::printf("everybody else left during the list: nothing to average with\n");
    } else {
        CHECK(rc);
        std::printf("list of %d averages done with %u peers, %llu bytes sent; recv[0][0] = %.3f (send %.3f)\n", kOps, total.local_world_size,
                    static_cast<unsigned long long>(total.tx_bytes), recv[0][0], send[0][0]);
    }
    std::fflush(stdout);
    CHECK(pcclDestroyCommunicator(comm));
    return 0;
}
