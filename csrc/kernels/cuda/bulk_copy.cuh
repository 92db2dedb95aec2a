// Tile movement with the bulk asynchronous copy engine (cp.async.bulk, SASS UBLKCP) for the HBM-bound phases of the
// collective kernels: staging an ordinary buffer into the arena and writing the arena out again.
//
// Why: a collective kernel runs ONE 512-thread block per SM (its blocks must be co-resident with their twins on the other
// GPUs), so a copy loop built from 16-byte loads keeps at most ~32 KB in flight per SM and needs registers for all of it;
// that reached 4.1-4.6 TB/s of the 6.5 TB/s the copy benchmark measures.  Here one thread per block drives a ring of shared
// memory stages: global -> shared with an mbarrier counting the bytes, shared -> global as bulk groups.  ~48 KB of loads
// and as many stores are in flight per SM without touching a register, and the other warps are free.
//
// Every copy here has BOTH ends in this GPU's own memory (the caller's buffer and this rank's own arena) and none is a
// multicast copy: peers' memory is only ever touched with ld/st/multimem instructions.  (TMA multicast loads here read only
// this GPU's memory, since on B200 a TMA multicast load from another GPU's memory hung the GPU.)
#pragma once
#include "collective.cuh"

namespace pccl::cuda {

constexpr int kBulkStages = 12;                 // ring of shared-memory stages
constexpr int kBulkAhead = 6;                   // loads in flight; the rest of the ring holds stores in flight
constexpr uint32_t kBulkTileBytes = kThreads * 16; // one tile of the work partition
constexpr size_t kBulkSmemBytes = size_t(kBulkStages) * kBulkTileBytes + kBulkStages * sizeof(uint64_t) + 128; // + alignment slack

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

struct BulkRing {
    uint8_t *stage;  // kBulkStages * kBulkTileBytes, 128-byte aligned
    uint64_t *mbar;  // kBulkStages
    uint32_t parity; // bit s = phase the next wait on stage s expects
    uint64_t loads, stores;
};

// All threads call once per kernel (before any copy); thread `driver` becomes the only user of the ring.
__device__ __forceinline__ void bulk_ring_init(BulkRing &r, uint8_t *smem, int driver) {
    uint8_t *base = reinterpret_cast<uint8_t *>((reinterpret_cast<uintptr_t>(smem) + 127) & ~uintptr_t(127));
    r.stage = base;
    r.mbar = reinterpret_cast<uint64_t *>(base + size_t(kBulkStages) * kBulkTileBytes);
    r.parity = 0;
    r.loads = r.stores = 0;
    if (static_cast<int>(threadIdx.x) == driver) {
        for (int s = 0; s < kBulkStages; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(r.mbar + s)));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("fence.proxy.async;" ::: "memory");
    }
    __syncthreads();
}

// driver thread only
__device__ __forceinline__ void bulk_store_oldest(BulkRing &r, const unsigned long long *dst_of, const uint32_t *bytes_of) {
    const int s = static_cast<int>(r.stores % kBulkStages);
    const uint32_t bar = smem_u32(r.mbar + s), want = (r.parity >> s) & 1u;
    uint32_t done = 0;
    while (!done) {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(done)
                     : "r"(bar), "r"(want)
                     : "memory");
    }
    r.parity ^= 1u << s;
    asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_of[s]), "r"(smem_u32(r.stage + size_t(s) * kBulkTileBytes)),
                 "r"(bytes_of[s])
                 : "memory");
    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    ++r.stores;
}

// Copies the tiles of packets [lo, hi) that this block owns from src to dst (both 16-byte aligned, both in THIS GPU's
// memory); packet v lives at byte v*16 of either.  Called by all threads of the participating group; only `driver` works.
// When it returns (to the driver) every byte has been written and is ordered before later generic-proxy accesses.
__device__ __forceinline__ void bulk_copy_owned_tiles(BulkRing &r, uint8_t *dst, const uint8_t *src, uint64_t lo, uint64_t hi, int driver) {
    if (static_cast<int>(threadIdx.x) != driver || lo >= hi) return;
    __shared__ uint32_t bytes_of[kBulkStages]; // per stage: length and destination of the tile it holds
    __shared__ unsigned long long dst_store[kBulkStages];
    asm volatile("fence.proxy.async;" ::: "memory"); // what this block synchronised on (flags, block barriers) precedes the async reads
    const uint64_t first_tile = lo / kThreads;
    uint64_t tile = first_tile + ((blockIdx.x + gridDim.x - first_tile % gridDim.x) % gridDim.x);
    for (; tile * kThreads < hi; tile += gridDim.x) {
        const uint64_t p0 = max(tile * uint64_t(kThreads), lo), p1 = min((tile + 1) * uint64_t(kThreads), hi);
        const uint32_t bytes = static_cast<uint32_t>((p1 - p0) * 16);
        // the stage this load takes was last used by the store issued kBulkStages loads ago
        if (r.loads >= kBulkStages) asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(kBulkStages - kBulkAhead - 1) : "memory");
        const int s = static_cast<int>(r.loads % kBulkStages);
        const uint32_t bar = smem_u32(r.mbar + s);
        bytes_of[s] = bytes;
        dst_store[s] = reinterpret_cast<unsigned long long>(dst + p0 * 16);
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                         smem_u32(r.stage + size_t(s) * kBulkTileBytes)),
                     "l"(src + p0 * 16), "r"(bytes), "r"(bar)
                     : "memory");
        ++r.loads;
        if (r.loads - r.stores > kBulkAhead) bulk_store_oldest(r, dst_store, bytes_of);
    }
    while (r.stores < r.loads) bulk_store_oldest(r, dst_store, bytes_of);
    asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    asm volatile("fence.proxy.async;" ::: "memory");
}

// host: dynamic shared memory for a launch that may use the ring (0 when it does not), with the opt-in done once per kernel
template <typename Kernel> inline size_t bulk_smem_for(Kernel kernel, const DevArgs &d) {
    if (!d.bulk) return 0;
    static thread_local const void *opted[64] = {};
    static thread_local int n_opted = 0;
    const void *fn = reinterpret_cast<const void *>(kernel);
    bool have = false;
    for (int i = 0; i < n_opted; ++i) have = have || opted[i] == fn;
    if (!have) {
        if (cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(kBulkSmemBytes)) != cudaSuccess) {
            (void) cudaGetLastError();
            return 0;
        }
        if (n_opted < 64) opted[n_opted++] = fn;
    }
    return kBulkSmemBytes;
}

} // namespace pccl::cuda

