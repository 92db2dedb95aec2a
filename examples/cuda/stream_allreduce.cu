// Device buffers, a CUDA stream, no host synchronisation in the loop: the extension API for peers on one NVLink domain.
//
//   ccoip_master &
//   ./stream_allreduce 127.0.0.1:48148 0     (GPU 0)
//   ./stream_allreduce 127.0.0.1:48148 1     (GPU 1)    ... one process per GPU
//
// Per step a kernel produces "gradients" into a buffer from the communicator's symmetric heap (peers reduce straight out of
// it: no staging copy), pcclExtAllReduceOnStream queues the averaging behind that kernel on the same stream, and a second
// kernel consumes the result -- still on the stream.  After the first operation on a tag the group holds a lease for it:
// later operations are launched without asking the master, so with enqueue=1 the call returns as soon as the kernel is
// queued and the host runs ahead.  Verdicts are collected by the next operation on the tag; pcclExtWaitStreamOps takes
// the last one.  A lost peer shows up there as pcclRankConnectionLost: pcclUpdateTopology, then carry on.
//
// Build:  nvcc -gencode arch=compute_100a,code=sm_100a -Iinclude examples/cuda/stream_allreduce.cu -Lpccl_b200/lib -lpccl \
//              -Xlinker -rpath,$PWD/pccl_b200/lib -o stream_allreduce
#include <pccl.h>

#include <cuda_runtime.h>

#include <cstdio>
#include <cstdlib>
#include <string>

#define CHECK(call)                                                                        \
    do {                                                                                   \
        const pcclResult_t rc_ = (call);                                                   \
        if (rc_ != pcclSuccess) {                                                          \
            std::fprintf(stderr, "%s failed with %d at line %d\n", #call, int(rc_), __LINE__); \
            std::exit(1);                                                                  \
        }                                                                                  \
    } while (0)
#define CUDA_CHECK(call)                                                                                   \
    do {                                                                                                   \
        const cudaError_t e_ = (call);                                                                     \
        if (e_ != cudaSuccess) {                                                                           \
            std::fprintf(stderr, "%s failed: %s at line %d\n", #call, cudaGetErrorString(e_), __LINE__);  \
            std::exit(1);                                                                                  \
        }                                                                                                  \
    } while (0)

__global__ void make_gradients(float *g, size_t n, int step, int rank_seed) {
    for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x)
        g[i] = float((i + step) % 5) + float(rank_seed);
}

__global__ void apply(float *weights, const float *g, size_t n, float lr) {
    for (size_t i = blockIdx.x * size_t(blockDim.x) + threadIdx.x; i < n; i += size_t(gridDim.x) * blockDim.x) weights[i] -= lr * g[i];
}

int main(int argc, char **argv) {
    const std::string master = argc > 1 ? argv[1] : "127.0.0.1:48148";
    const int device = argc > 2 ? std::atoi(argv[2]) : 0;
    const int steps = argc > 3 ? std::atoi(argv[3]) : 50;
    const int want_peers = argc > 4 ? std::atoi(argv[4]) : 2;
    const size_t n = size_t(32) << 20; // 128 MiB of float32
    unsigned b[4], port = CCOIP_PROTOCOL_PORT_MASTER;
    if (std::sscanf(master.c_str(), "%u.%u.%u.%u:%u", &b[0], &b[1], &b[2], &b[3], &port) < 4) return 2;

    CUDA_CHECK(cudaSetDevice(device)); // before the communicator: its arena lives on the current device
    CHECK(pcclInit());
    pcclCommCreateParams_t params{};
    params.master_address.inet.protocol = inetIPv4;
    for (int i = 0; i < 4; ++i) params.master_address.inet.ipv4.data[i] = uint8_t(b[i]);
    params.master_address.port = uint16_t(port);
    pcclComm_t *comm = nullptr;
    CHECK(pcclCreateCommunicator(&params, &comm));
    CHECK(pcclConnect(comm));
    int world = 0;
    for (;;) {
        CHECK(pcclGetAttribute(comm, PCCL_ATTRIBUTE_GLOBAL_WORLD_SIZE, &world));
        if (world >= want_peers) break;
        CHECK(pcclUpdateTopology(comm));
    }

    float *gradients = nullptr, *weights = nullptr;
    CHECK(pcclExtSymmetricAlloc(comm, n * sizeof(float), reinterpret_cast<void **>(&gradients)));
    CUDA_CHECK(cudaMalloc(&weights, n * sizeof(float)));
    CUDA_CHECK(cudaMemset(weights, 0, n * sizeof(float)));
    cudaStream_t stream;
    CUDA_CHECK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    cudaEvent_t t0, t1;
    CUDA_CHECK(cudaEventCreate(&t0));
    CUDA_CHECK(cudaEventCreate(&t1));

    pcclReduceDescriptor_t desc{};
    desc.count = n, desc.op = pcclAvg, desc.tag = 1;
    desc.src_descriptor.datatype = pcclFloat;
    for (int step = 0; step < steps; ++step) {
        if (step == 1) CUDA_CHECK(cudaEventRecord(t0, stream)); // step 0 votes at the master and obtains the lease
        make_gradients<<<296, 512, 0, stream>>>(gradients, n, step, device);
        pcclReduceInfo_t info{};
        const pcclResult_t rc = pcclExtAllReduceOnStream(gradients, gradients, &desc, comm, stream, /*enqueue=*/1, &info);
        if (rc == pcclRankConnectionLost) { // a peer left: settle what is in flight, re-form the group, repeat the step
            pcclExtWaitStreamOps(comm, 0, /*all_tags=*/1, nullptr);
            CHECK(pcclUpdateTopology(comm));
            --step;
            continue;
        }
        CHECK(rc);
        apply<<<296, 512, 0, stream>>>(weights, gradients, n, 0.01f);
    }
    CUDA_CHECK(cudaEventRecord(t1, stream));
    pcclReduceInfo_t last{};
    CHECK(pcclExtWaitStreamOps(comm, 1, 0, &last)); // verdict of the last enqueued operation
    CUDA_CHECK(cudaStreamSynchronize(stream));
    float ms = 0;
    CUDA_CHECK(cudaEventElapsedTime(&ms, t0, t1));
    pcclExtReduceStats_t stats{};
    CHECK(pcclExtGetLastReduceStats(comm, 1, &stats));
    float first[4];
    CUDA_CHECK(cudaMemcpy(first, weights, sizeof first, cudaMemcpyDeviceToHost));
    std::printf("gpu %d: %d steps with %u peers, %.3f ms per step (gradients + all-reduce of %zu MiB + update), plane %d, leased %u, weights[0..3] = %.3f %.3f %.3f %.3f\n",
                device, steps, last.local_world_size, steps > 1 ? ms / float(steps - 1) : 0.f, n * sizeof(float) >> 20, int(stats.plane), stats.leased, first[0],
                first[1], first[2], first[3]);
    CHECK(pcclExtSymmetricFree(comm, gradients));
    CUDA_CHECK(cudaFree(weights));
    CHECK(pcclDestroyCommunicator(comm));
    return 0;
}
