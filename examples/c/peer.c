/*
 * A complete fault-tolerant peer in C99: the loop every PCCL application is built around.
 *
 *   ./ccoip_master &                      (pccl_b200/bin/ccoip_master, listens on 0.0.0.0:48148)
 *   ./peer 127.0.0.1 [steps] [min_peers] [ms of pretended work per step]
 *                                         (start as many as you like, at any time; kill any of them at any time)
 *
 * Per step:  admit newcomers -> agree on the shared state (weights + step counter) -> "train" -> average the
 * "gradients" -> apply them.  A failed collective is retried with whoever is left; a peer that finds itself alone waits
 * for company.  Counterpart of the reference's docs/md/02-Getting Started/03_Basic_HelloWorld.md and
 * tests/basic_reduce_test/main.cpp, written against include/pccl.h only.
 *
 * Build:  gcc -std=c99 -Iinclude examples/c/peer.c -Lpccl_b200/lib -lpccl -Wl,-rpath,$PWD/pccl_b200/lib -o peer
 */
#define _POSIX_C_SOURCE 199309L /* nanosleep */
#include <pccl.h>

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define N_WEIGHTS 4096

#define CHECK(call)                                                                     \
    do {                                                                                \
        pcclResult_t rc_ = (call);                                                      \
        if (rc_ != pcclSuccess) {                                                       \
            fprintf(stderr, "%s failed with %d at line %d\n", #call, (int) rc_, __LINE__); \
            exit(1);                                                                    \
        }                                                                               \
    } while (0)

static int world_size(const pcclComm_t *comm, pcclAttribute_t which) {
    int value = 0;
    CHECK(pcclGetAttribute(comm, which, &value));
    return value;
}

int main(int argc, char **argv) {
    const char *ip = argc > 1 ? argv[1] : "127.0.0.1";
    const long steps = argc > 2 ? atol(argv[2]) : 20;
    const int min_peers = argc > 3 ? atoi(argv[3]) : 2;
    const long work_ms = argc > 4 ? atol(argv[4]) : 0;
    unsigned a, b, c, d;
    if (sscanf(ip, "%u.%u.%u.%u", &a, &b, &c, &d) != 4) {
        fprintf(stderr, "usage: %s <master ipv4> [steps] [min_peers] [work_ms]\n", argv[0]);
        return 2;
    }

    CHECK(pcclInit());

    pcclCommCreateParams_t params;
    memset(&params, 0, sizeof params); /* zero = defaults: group 0, one connection per neighbour, standard ports */
    params.master_address.inet.protocol = inetIPv4;
    params.master_address.inet.ipv4.data[0] = (uint8_t) a;
    params.master_address.inet.ipv4.data[1] = (uint8_t) b;
    params.master_address.inet.ipv4.data[2] = (uint8_t) c;
    params.master_address.inet.ipv4.data[3] = (uint8_t) d;
    params.master_address.port = CCOIP_PROTOCOL_PORT_MASTER;
    if (getenv("PCCL_EXAMPLE_MASTER_PORT")) params.master_address.port = (uint16_t) atoi(getenv("PCCL_EXAMPLE_MASTER_PORT"));

    pcclComm_t *comm = NULL;
    CHECK(pcclCreateCommunicator(&params, &comm));
    CHECK(pcclConnect(comm)); /* returns once the peers already in the run have admitted this one */

    static float weights[N_WEIGHTS], gradients[N_WEIGHTS];
    uint64_t step = 0; /* travels with the weights: a late joiner continues where the run is */
    pcclTensorInfo_t infos[2];
    memset(infos, 0, sizeof infos);
    infos[0].name = "weights", infos[0].data = weights, infos[0].count = N_WEIGHTS, infos[0].datatype = pcclFloat, infos[0].device_type = pcclDeviceCpu;
    infos[1].name = "step", infos[1].data = &step, infos[1].count = 1, infos[1].datatype = pcclUint64, infos[1].device_type = pcclDeviceCpu;
    pcclSharedState_t state;
    state.revision = 0, state.count = 2, state.infos = infos;

    int first = 1;
    while ((long) step < steps) {
        /* 1. membership.  Not in the very first iteration: the peers that admitted us are already past this call and
         *    are waiting for us in the next one. */
        if (!first) {
            bool pending = false;
            CHECK(pcclArePeersPending(comm, &pending));
            if (pending || world_size(comm, PCCL_ATTRIBUTE_GLOBAL_WORLD_SIZE) < min_peers) {
                pcclResult_t rc = pcclUpdateTopology(comm);
                if (rc != pcclSuccess && rc != pcclUpdateTopologyFailed) CHECK(rc); /* a failed round is simply tried again */
            }
        }
        first = 0;
        if (world_size(comm, PCCL_ATTRIBUTE_PEER_GROUP_WORLD_SIZE) < min_peers) continue; /* alone: wait for company */

        /* 2. shared state: whoever differs from the majority (a newcomer, a peer that missed an update) receives */
        pcclSharedStateSyncInfo_t sync_info;
        pcclResult_t rc = pcclSynchronizeSharedState(comm, &state, PCCL_SHARED_STATE_SYNC_STRATEGY_ENFORCE_POPULAR, &sync_info);
        if (rc != pcclSuccess) continue; /* somebody left in the middle: next round */
        if (sync_info.rx_bytes) printf("received %llu bytes of state, continuing at step %llu\n", (unsigned long long) sync_info.rx_bytes, (unsigned long long) step);

        /* 3. local work */
        for (int i = 0; i < N_WEIGHTS; ++i) gradients[i] = (float) ((i + step) % 7) - 3.0f;
        if (work_ms > 0) {
            struct timespec ts;
            ts.tv_sec = work_ms / 1000, ts.tv_nsec = (work_ms % 1000) * 1000000L;
            nanosleep(&ts, NULL);
        }

        /* 4. average; in place.  On failure the input is intact and the group has shrunk: retry. */
        pcclReduceDescriptor_t desc;
        memset(&desc, 0, sizeof desc);
        desc.count = N_WEIGHTS, desc.op = pcclAvg, desc.tag = 0;
        desc.src_descriptor.datatype = pcclFloat;
        pcclReduceInfo_t info;
        int reduced = 0;
        while (world_size(comm, PCCL_ATTRIBUTE_PEER_GROUP_WORLD_SIZE) >= 2) {
            rc = pcclAllReduce(gradients, gradients, &desc, comm, &info);
            if (rc == pcclSuccess) {
                reduced = 1;
                break;
            }
            if (rc != pcclRankConnectionLost) CHECK(rc);
        }
        if (!reduced) continue; /* everybody else left: the step is dropped and repeated once there is company again */

        /* 5. every peer applies the same update, so the shared state stays identical without moving a byte */
        for (int i = 0; i < N_WEIGHTS; ++i) weights[i] -= 0.01f * gradients[i];
        ++step;
        ++state.revision; /* exactly one per synchronisation */
        printf("step %llu done with %u peers (sent %llu bytes)\n", (unsigned long long) step, info.local_world_size, (unsigned long long) info.tx_bytes);
        fflush(stdout);
    }
    printf("finished: weights[0..2] = %.4f %.4f %.4f\n", weights[0], weights[1], weights[2]);
    CHECK(pcclDestroyCommunicator(comm));
    return 0;
}
