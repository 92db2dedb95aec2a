// All-reduce between GPUs that can address each other's memory (NVLink 5 / NVSwitch).
//
// There is no counterpart in the reference: its all-reduce is a host-memory TCP ring
// (ccoip/src/cpp/reduce.cpp:528-784, element kernels reduce_kernels.cpp:56-215, AVG :437-510, min-max
// quantiser quantize_kernels.cpp:24-217).  This file replaces that whole path for peers in one
// NVLink domain with a single kernel launch per rank that fuses staging, reduction, AVG
// finalisation and (optionally) block-scaled quantise -> reduce -> dequantise.
//
// Memory model.  Every rank owns an "arena" that all ranks have mapped.  It starts with a
// control page of barrier flags, staging data follows at `data_offset`.
//   flags[slot][block][src_rank] : u64, written with st.release.sys by block `block` of rank
//   `src_rank`, polled with ld.acquire.sys by the same block of the owner.
// Flag values only grow (the caller derives them from the master's sequence number), so values
// left behind by an aborted operation can never satisfy a later wait.
//
// Work partition.  Data is cut into tiles of THREADS 16-byte packets.  Tile t belongs to block
// t % gridDim.x ON EVERY RANK AND IN EVERY PHASE.  A block therefore only ever consumes data that
// the same-numbered block of a peer produced, which is what makes per-block barriers sufficient
// (no grid-wide synchronisation, no cooperative launch) and lets blocks run ahead of each other.
//
// Two-shot algorithm (bandwidth regime):
//   phase 0  copy own input into own arena                      (skipped for symmetric memory)
//   barrier
//   phase 1  rank r reduces chunk r: reads chunk r of every peer's arena, reduces in rank order,
//            applies AVG, writes the result to its arena AND to its own output
//   barrier
//   phase 2  reads the reduced chunks of all other ranks from their arenas into the output
//   barrier  (nobody may overwrite an arena a peer is still reading)
// One-shot algorithm (latency regime): phase 0, barrier, every rank reduces everything, barrier.
//
// Every spin is bounded: it polls a host-mapped abort word (set by the control thread when the
// master aborts the operation or a peer dies) and a cycle budget.
#include <cstdio>

#include "common.cuh"
#include "outer_step.cuh"
#include "pccl_kernels.h"
#include "quant.cuh"

namespace pccl::cuda {

constexpr int kThreads = 512;
constexpr size_t kFlagBankBytes = size_t(PCCL_K_MAX_BLOCKS) * 128; // one 128-byte line per block
constexpr int kFlagBanks = 16;
constexpr size_t kControlBytes = 1u << 20;
static_assert(kFlagBanks * kFlagBankBytes <= kControlBytes, "control page too small");
static_assert(PCCL_K_MAX_RANKS * 8 <= 128, "flag line too small");

struct DevArgs {
    uint8_t *arena[PCCL_K_MAX_RANKS];
    uint32_t rank, world, flag_slot;
    uint64_t flag_base, data_offset;
    const void *src;
    void *dst;
    uint64_t count;
    uint64_t avg_divisor; // 0: none
    int src_in_arena;
    volatile uint32_t *abort_flag;
    uint32_t *status;
    uint64_t timeout_cycles;
};

// Peer base pointers live in shared memory: they are indexed with run-time rank numbers, and
// a run-time index into kernel parameters would be demoted to a per-thread stack copy.
struct PeerTable {
    uint8_t *base[PCCL_K_MAX_RANKS];
};
__device__ __forceinline__ void load_peer_table(PeerTable &t, const DevArgs &a) {
    if (threadIdx.x < PCCL_K_MAX_RANKS) t.base[threadIdx.x] = threadIdx.x < a.world ? a.arena[threadIdx.x] + a.data_offset : nullptr;
    __syncthreads();
}
constexpr int kPeerGroup = 8; // peers whose loads are kept in flight together

__device__ __forceinline__ uint64_t *flag_ptr(const DevArgs &a, uint8_t *arena_base, uint32_t block, uint32_t writer) {
    return reinterpret_cast<uint64_t *>(arena_base + a.flag_slot * kFlagBankBytes + size_t(block) * 128) + writer;
}

// Cross-GPU barrier among the same-numbered blocks of all ranks.  Returns false when the
// operation was aborted or timed out (uniform across the block).
__device__ bool peer_barrier(const DevArgs &a, const PeerTable &t, uint64_t value, bool record_failure = true) {
    __shared__ int s_fail;
    if (threadIdx.x == 0) s_fail = 0;
    __syncthreads(); // this block's earlier writes are ordered before the flag stores below
    if (threadIdx.x < a.world) {
        const uint32_t peer = threadIdx.x;
        st_flag_release(flag_ptr(a, t.base[peer] - a.data_offset, blockIdx.x, a.rank), value);
        const uint64_t *mine = flag_ptr(a, t.base[a.rank] - a.data_offset, blockIdx.x, peer);
        const long long t0 = clock64();
        uint32_t spins = 0;
        while (ld_flag_acquire(mine) < value) {
            if ((++spins & 0x3ffu) == 0) {
                if (*a.abort_flag != 0) {
                    atomicExch(&s_fail, 1);
                    break;
                }
                if (static_cast<uint64_t>(clock64() - t0) > a.timeout_cycles) {
                    atomicExch(&s_fail, 2);
                    break;
                }
            }
        }
    }
    __syncthreads();
    const int fail = s_fail;
    if (fail != 0 && record_failure && threadIdx.x == 0) atomicMax(a.status, static_cast<uint32_t>(fail));
    return fail == 0;
}

// Iterates over the packets of [lo, hi) that belong to this block, U tiles at a time.
// fn(v[], n_valid) gets up to U packet indices (all < hi).
template <int U, typename Fn> __device__ __forceinline__ void for_owned_packets(uint64_t lo, uint64_t hi, Fn fn) {
    if (lo >= hi) return;
    const uint64_t first_tile = lo / kThreads;
    // smallest tile >= first_tile owned by this block
    uint64_t tile = first_tile + ((blockIdx.x + gridDim.x - first_tile % gridDim.x) % gridDim.x);
    const uint64_t stride = uint64_t(gridDim.x) * U;
    for (; tile * kThreads < hi; tile += stride) {
        uint64_t v[U];
        bool ok[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            v[u] = (tile + uint64_t(u) * gridDim.x) * kThreads + threadIdx.x;
            ok[u] = v[u] >= lo && v[u] < hi;
        }
        fn(v, ok);
    }
}

template <typename T> __device__ __forceinline__ uint4 load_input(const T *src, uint64_t v, uint64_t count, bool aligned) {
    constexpr int N = 16 / sizeof(T);
    const uint64_t e = v * N;
    if (aligned && e + N <= count) return ld_stream_16(src + e);
    Pack<T> p;
#pragma unroll
    for (int i = 0; i < N; ++i) p.v[i] = e + i < count ? src[e + i] : T{};
    return *reinterpret_cast<uint4 *>(&p);
}

template <typename T> __device__ __forceinline__ void store_output(T *dst, uint64_t v, uint64_t count, bool aligned, uint4 val) {
    constexpr int N = 16 / sizeof(T);
    const uint64_t e = v * N;
    if (aligned && e + N <= count) {
        st_stream_16(dst + e, val);
        return;
    }
    const Pack<T> p = *reinterpret_cast<Pack<T> *>(&val);
#pragma unroll
    for (int i = 0; i < N; ++i)
        if (e + i < count) dst[e + i] = p.v[i];
}

template <typename T, int OP> __device__ __forceinline__ void accumulate(typename Elem<T>::acc_t (&acc)[16 / sizeof(T)], uint4 raw, bool first) {
    constexpr int N = 16 / sizeof(T);
    const Pack<T> p = *reinterpret_cast<const Pack<T> *>(&raw);
#pragma unroll
    for (int i = 0; i < N; ++i) {
        const auto x = Elem<T>::load(p.v[i]);
        acc[i] = first ? x : combine<OP>(acc[i], x);
    }
}

template <typename T> __device__ __forceinline__ uint4 finish(typename Elem<T>::acc_t (&acc)[16 / sizeof(T)], uint64_t avg_divisor) {
    constexpr int N = 16 / sizeof(T);
    using A = typename Elem<T>::acc_t;
    Pack<T> p;
#pragma unroll
    for (int i = 0; i < N; ++i) {
        A x = acc[i];
        if (avg_divisor) x = x / static_cast<A>(avg_divisor);
        p.v[i] = Elem<T>::store(x);
    }
    return *reinterpret_cast<uint4 *>(&p);
}

// -----------------------------------------------------------------------------------------
// two-shot, unquantised
// -----------------------------------------------------------------------------------------
template <typename T, int OP> __global__ void __launch_bounds__(kThreads, 1) allreduce_two_shot_kernel(const DevArgs a) {
    constexpr int N = 16 / sizeof(T);
    __shared__ PeerTable tbl;
    load_peer_table(tbl, a);
    const uint64_t n_pack = (a.count + N - 1) / N;
    const uint64_t per = (n_pack + a.world - 1) / a.world; // packets per chunk
    const T *src = static_cast<const T *>(a.src);
    T *dst = static_cast<T *>(a.dst);
    const bool src_aligned = (reinterpret_cast<uintptr_t>(src) & 15) == 0, dst_aligned = (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
    uint4 *mine = reinterpret_cast<uint4 *>(tbl.base[a.rank]);

    // phase 0: stage the input where the peers can read it
    if (!a.src_in_arena) {
        for_owned_packets<4>(0, n_pack, [&](const uint64_t(&v)[4], const bool(&ok)[4]) {
            uint4 tmp[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u]) tmp[u] = load_input(src, v[u], a.count, src_aligned);
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u]) mine[v[u]] = tmp[u];
        });
    }
    if (!peer_barrier(a, tbl, a.flag_base + 1)) return;

    // phase 1: reduce the chunk this rank owns, peers in rank order (deterministic result)
    {
        const uint64_t lo = min(uint64_t(a.rank) * per, n_pack), hi = min(lo + per, n_pack);
        for_owned_packets<2>(lo, hi, [&](const uint64_t(&v)[2], const bool(&ok)[2]) {
            typename Elem<T>::acc_t acc[2][N];
            for (uint32_t r0 = 0; r0 < a.world; r0 += kPeerGroup) {
                uint4 raw[2][kPeerGroup];
#pragma unroll
                for (int u = 0; u < 2; ++u)
#pragma unroll
                    for (int g = 0; g < kPeerGroup; ++g)
                        if (r0 + g < a.world && ok[u]) raw[u][g] = ld_peer_16(reinterpret_cast<const uint4 *>(tbl.base[r0 + g]) + v[u]);
#pragma unroll
                for (int u = 0; u < 2; ++u)
#pragma unroll
                    for (int g = 0; g < kPeerGroup; ++g)
                        if (r0 + g < a.world && ok[u]) accumulate<T, OP>(acc[u], raw[u][g], r0 + g == 0);
            }
#pragma unroll
            for (int u = 0; u < 2; ++u) {
                if (!ok[u]) continue;
                const uint4 out = finish<T>(acc[u], a.avg_divisor);
                mine[v[u]] = out;
                store_output(dst, v[u], a.count, dst_aligned, out);
            }
        });
    }
    if (!peer_barrier(a, tbl, a.flag_base + 2)) return;

    // phase 2: collect the other ranks' reduced chunks, starting with the right-hand neighbour
    // so that at any moment the ranks pull from different peers
    for (uint32_t k = 1; k < a.world; ++k) {
        const uint32_t peer = (a.rank + k) % a.world;
        const uint4 *from = reinterpret_cast<const uint4 *>(tbl.base[peer]);
        const uint64_t lo = min(uint64_t(peer) * per, n_pack), hi = min(lo + per, n_pack);
        for_owned_packets<4>(lo, hi, [&](const uint64_t(&v)[4], const bool(&ok)[4]) {
            uint4 tmp[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u]) tmp[u] = ld_peer_16(from + v[u]);
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u]) store_output(dst, v[u], a.count, dst_aligned, tmp[u]);
        });
    }
    peer_barrier(a, tbl, a.flag_base + 3);
}

// -----------------------------------------------------------------------------------------
// one-shot, unquantised
// -----------------------------------------------------------------------------------------
template <typename T, int OP> __global__ void __launch_bounds__(kThreads, 1) allreduce_one_shot_kernel(const DevArgs a) {
    constexpr int N = 16 / sizeof(T);
    __shared__ PeerTable tbl;
    load_peer_table(tbl, a);
    const uint64_t n_pack = (a.count + N - 1) / N;
    const T *src = static_cast<const T *>(a.src);
    T *dst = static_cast<T *>(a.dst);
    const bool src_aligned = (reinterpret_cast<uintptr_t>(src) & 15) == 0, dst_aligned = (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
    uint4 *mine = reinterpret_cast<uint4 *>(tbl.base[a.rank]);
    if (!a.src_in_arena) {
        for_owned_packets<1>(0, n_pack, [&](const uint64_t(&v)[1], const bool(&ok)[1]) {
            if (ok[0]) mine[v[0]] = load_input(src, v[0], a.count, src_aligned);
        });
    }
    if (!peer_barrier(a, tbl, a.flag_base + 1)) return;
    for_owned_packets<1>(0, n_pack, [&](const uint64_t(&v)[1], const bool(&ok)[1]) {
        if (!ok[0]) return;
        typename Elem<T>::acc_t acc[N];
        for (uint32_t r0 = 0; r0 < a.world; r0 += kPeerGroup) { // fixed order: bit-identical on every rank
            uint4 raw[kPeerGroup];
#pragma unroll
            for (int g = 0; g < kPeerGroup; ++g)
                if (r0 + g < a.world) raw[g] = ld_peer_16(reinterpret_cast<const uint4 *>(tbl.base[r0 + g]) + v[0]);
#pragma unroll
            for (int g = 0; g < kPeerGroup; ++g)
                if (r0 + g < a.world) accumulate<T, OP>(acc, raw[g], r0 + g == 0);
        }
        store_output(dst, v[0], a.count, dst_aligned, finish<T>(acc, a.avg_divisor));
    });
    peer_barrier(a, tbl, a.flag_base + 2);
}

// -----------------------------------------------------------------------------------------
// two-shot with block-scaled wire format (fp32 in memory, 1 byte per element on NVLink)
//   staging layout: [u32 codes, one per float4][block records, one per 32 float4]
// -----------------------------------------------------------------------------------------
template <int ALGO, int OP> __global__ void __launch_bounds__(kThreads, 1) allreduce_two_shot_quant_kernel(const DevArgs a) {
    // Work unit = one 16-byte code packet = 16 fp32 values of one thread; 8 lanes share a block record.
    constexpr int MB = QuantMeta<ALGO>::kBytes;
    const uint64_t n_blocks = (a.count + PCCL_K_QUANT_BLOCK - 1) / PCCL_K_QUANT_BLOCK;
    const uint64_t n_pack = n_blocks * 8;                                     // code packets incl. padding
    const uint64_t per = ((n_blocks + a.world - 1) / a.world) * 8;            // chunk = whole blocks
    const float *src = static_cast<const float *>(a.src);
    float *dst = static_cast<float *>(a.dst);
    const bool src_aligned = (reinterpret_cast<uintptr_t>(src) & 15) == 0, dst_aligned = (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
    const int lane = threadIdx.x & 31;
    __shared__ PeerTable tbl;
    load_peer_table(tbl, a);
    auto codes = [&](uint32_t r) { return reinterpret_cast<uint4 *>(tbl.base[r]); };
    auto metas = [&](uint32_t r) { return tbl.base[r] + n_pack * 16; };

    auto valid_of = [&](uint64_t v) -> int {
        const uint64_t e = v * 16;
        return e >= a.count ? 0 : (a.count - e >= 16 ? 16 : static_cast<int>(a.count - e));
    };
    // The input is read with plain (L1-allocating) loads: a thread walks 64 contiguous bytes in
    // four steps, the neighbouring lanes fill the rest of each 128-byte line.
    auto load16 = [&](uint64_t v, float (&x)[16]) -> int {
        const int valid = valid_of(v);
        if (valid == 16 && src_aligned) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float4 f = *reinterpret_cast<const float4 *>(src + v * 16 + j * 4);
                x[4 * j] = f.x, x[4 * j + 1] = f.y, x[4 * j + 2] = f.z, x[4 * j + 3] = f.w;
            }
        } else {
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] = i < valid ? src[v * 16 + i] : 0.0f;
        }
        return valid;
    };
    auto store16 = [&](uint64_t v, const float (&y)[16]) {
        const int valid = valid_of(v);
        if (valid == 16 && dst_aligned) {
#pragma unroll
            for (int j = 0; j < 4; ++j) *reinterpret_cast<float4 *>(dst + v * 16 + j * 4) = make_float4(y[4 * j], y[4 * j + 1], y[4 * j + 2], y[4 * j + 3]);
        } else {
#pragma unroll
            for (int i = 0; i < 16; ++i)
                if (i < valid) dst[v * 16 + i] = y[i];
        }
    };
    auto store_record = [&](uint32_t r, uint64_t v, uint32_t lo, uint32_t hi) {
        if ((lane & 7) != 0) return;
        uint8_t *m = metas(r) + (v / 8) * MB;
        if constexpr (MB == 8)
            *reinterpret_cast<uint2 *>(m) = make_uint2(lo, hi);
        else
            *reinterpret_cast<uint32_t *>(m) = lo;
    };
    auto load_record = [&](uint32_t r, uint64_t v, uint32_t &lo, uint32_t &hi) {
        const uint8_t *m = metas(r) + (v / 8) * MB;
        if constexpr (MB == 8) {
            const uint2 q = ld_peer_8(m);
            lo = q.x, hi = q.y;
        } else {
            lo = ld_peer_4(m), hi = 0;
        }
    };

    // Ranges are multiples of 8 packets and tiles of 512, so `ok` is uniform over every 8-lane
    // group (NOT over the warp): the shuffles inside quantize16 use a per-group mask.
    // phase 0: quantise own input into own arena
    for_owned_packets<1>(0, n_pack, [&](const uint64_t(&v)[1], const bool(&ok)[1]) {
        if (!ok[0]) return;
        float x[16];
        const int valid = load16(v[0], x);
        uint32_t lo, hi;
        const uint4 packed = quantize16<ALGO>(x, valid, lo, hi);
        codes(a.rank)[v[0]] = packed;
        store_record(a.rank, v[0], lo, hi);
    });
    if (!peer_barrier(a, tbl, a.flag_base + 1)) return;

    // phase 1: dequantise + reduce own chunk in fp32, re-quantise once, adopt the decoded value
    {
        const uint64_t lo_p = min(uint64_t(a.rank) * per, n_pack), hi_p = min(lo_p + per, n_pack);
        for_owned_packets<1>(lo_p, hi_p, [&](const uint64_t(&v)[1], const bool(&ok)[1]) {
            if (!ok[0]) return;
            float acc[16];
            for (uint32_t r0 = 0; r0 < a.world; r0 += kPeerGroup) {
                uint4 raw[kPeerGroup];
                uint32_t mlo[kPeerGroup], mhi[kPeerGroup];
#pragma unroll
                for (int g = 0; g < kPeerGroup; ++g)
                    if (r0 + g < a.world) {
                        raw[g] = ld_peer_16(codes(r0 + g) + v[0]);
                        load_record(r0 + g, v[0], mlo[g], mhi[g]);
                    }
#pragma unroll
                for (int g = 0; g < kPeerGroup; ++g)
                    if (r0 + g < a.world) {
                        float x[16];
                        dequantize16<ALGO>(raw[g], mlo[g], mhi[g], x);
#pragma unroll
                        for (int i = 0; i < 16; ++i) acc[i] = r0 + g == 0 ? x[i] : combine<OP>(acc[i], x[i]);
                    }
            }
            if (a.avg_divisor)
#pragma unroll
                for (int i = 0; i < 16; ++i) acc[i] = acc[i] / static_cast<float>(a.avg_divisor);
            uint32_t lo, hi;
            const uint4 packed = quantize16<ALGO>(acc, valid_of(v[0]), lo, hi);
            codes(a.rank)[v[0]] = packed;
            store_record(a.rank, v[0], lo, hi);
            float y[16]; // the owner keeps exactly what the others will decode
            dequantize16<ALGO>(packed, lo, hi, y);
            store16(v[0], y);
        });
    }
    if (!peer_barrier(a, tbl, a.flag_base + 2)) return;

    // phase 2: fetch the other chunks as codes, decode into the output
    for (uint32_t k = 1; k < a.world; ++k) {
        const uint32_t peer = (a.rank + k) % a.world;
        const uint64_t lo_p = min(uint64_t(peer) * per, n_pack), hi_p = min(lo_p + per, n_pack);
        for_owned_packets<2>(lo_p, hi_p, [&](const uint64_t(&v)[2], const bool(&ok)[2]) {
            uint4 raw[2];
            uint32_t mlo[2], mhi[2];
#pragma unroll
            for (int u = 0; u < 2; ++u)
                if (ok[u]) {
                    raw[u] = ld_peer_16(codes(peer) + v[u]);
                    load_record(peer, v[u], mlo[u], mhi[u]);
                }
#pragma unroll
            for (int u = 0; u < 2; ++u)
                if (ok[u]) {
                    float y[16];
                    dequantize16<ALGO>(raw[u], mlo[u], mhi[u], y);
                    store16(v[u], y);
                }
        });
    }
    peer_barrier(a, tbl, a.flag_base + 3);
}

// -----------------------------------------------------------------------------------------
// DiLoCo outer step in one launch (fp32, AVG), peer-to-peer flavour of outer_step_nvls_kernel:
//   phase 0  arena = params - inner
//   phase 1  owner of a chunk averages it over all arenas, result goes to its arena ONLY
//   phase 2  every rank walks all chunks (its own included) and applies the Nesterov update
// The caller's tensors are first written in phase 2, when every rank provably holds the whole
// average; a failure of the closing barrier (a peer died after that point) is therefore not an
// error of this operation and is not recorded.
// -----------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads, 1) outer_step_two_shot_kernel(const DevArgs a, const OuterStepDev s) {
    __shared__ PeerTable tbl;
    load_peer_table(tbl, a);
    const uint64_t n_pack = (a.count + 3) / 4;
    const uint64_t per = (n_pack + a.world - 1) / a.world;
    uint4 *mine = reinterpret_cast<uint4 *>(tbl.base[a.rank]);

    for_owned_packets<4>(0, n_pack, [&](const uint64_t(&v)[4], const bool(&ok)[4]) {
        uint4 d[4];
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (ok[u]) d[u] = outer_delta_packet(s, v[u], a.count);
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (ok[u]) mine[v[u]] = d[u];
    });
    if (!peer_barrier(a, tbl, a.flag_base + 1)) return;

    {
        const uint64_t lo = min(uint64_t(a.rank) * per, n_pack), hi = min(lo + per, n_pack);
        for_owned_packets<2>(lo, hi, [&](const uint64_t(&v)[2], const bool(&ok)[2]) {
            float acc[2][4];
            for (uint32_t r0 = 0; r0 < a.world; r0 += kPeerGroup) {
                uint4 raw[2][kPeerGroup];
#pragma unroll
                for (int u = 0; u < 2; ++u)
#pragma unroll
                    for (int g = 0; g < kPeerGroup; ++g)
                        if (r0 + g < a.world && ok[u]) raw[u][g] = ld_peer_16(reinterpret_cast<const uint4 *>(tbl.base[r0 + g]) + v[u]);
#pragma unroll
                for (int u = 0; u < 2; ++u)
#pragma unroll
                    for (int g = 0; g < kPeerGroup; ++g)
                        if (r0 + g < a.world && ok[u]) accumulate<float, OP_SUM>(acc[u], raw[u][g], r0 + g == 0);
            }
#pragma unroll
            for (int u = 0; u < 2; ++u)
                if (ok[u]) mine[v[u]] = finish<float>(acc[u], a.avg_divisor);
        });
    }
    if (!peer_barrier(a, tbl, a.flag_base + 2)) return;

    for (uint32_t k = 0; k < a.world; ++k) {
        const uint32_t peer = (a.rank + k) % a.world;
        const uint4 *from = reinterpret_cast<const uint4 *>(tbl.base[peer]);
        const uint64_t lo = min(uint64_t(peer) * per, n_pack), hi = min(lo + per, n_pack);
        for_owned_packets<4>(lo, hi, [&](const uint64_t(&v)[4], const bool(&ok)[4]) {
            uint4 g[4];
            OuterLoaded l[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u]) {
                    g[u] = ld_peer_16(from + v[u]);
                    l[u] = outer_load_packet(s, v[u], a.count);
                }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u]) outer_apply_packet(s, v[u], a.count, g[u], l[u]);
        });
    }
    peer_barrier(a, tbl, a.flag_base + 3, false);
}

} // namespace pccl::cuda

using namespace pccl::cuda;

extern "C" size_t pcclKernelP2PControlBytes(void) { return kControlBytes; }

extern "C" size_t pcclKernelP2PStagingBytes(uint64_t count, int dtype, int quant_algorithm) {
    if (quant_algorithm == QUANT_BLOCK_MINMAX || quant_algorithm == QUANT_BLOCK_FP8) {
        const uint64_t n_blocks = (count + PCCL_K_QUANT_BLOCK - 1) / PCCL_K_QUANT_BLOCK;
        const uint64_t mb = quant_algorithm == QUANT_BLOCK_MINMAX ? 8 : 4;
        return n_blocks * PCCL_K_QUANT_BLOCK + n_blocks * mb + 256;
    }
    const size_t esz = pcclExtDataTypeSize(static_cast<pcclDataType_t>(dtype));
    return (count * esz + 15) / 16 * 16 + 256;
}

extern "C" int pcclKernelAllReduceP2P(const pcclKernelP2PArgs_t *args, int *algorithm_out, void *stream_v) {
    if (!args || args->world < 1 || args->world > PCCL_K_MAX_RANKS || args->rank >= args->world) return cudaErrorInvalidValue;
    if (args->flag_slot >= static_cast<uint32_t>(kFlagBanks)) return cudaErrorInvalidValue;
    const auto stream = static_cast<cudaStream_t>(stream_v);
    DevArgs d{};
    for (uint32_t r = 0; r < args->world; ++r) d.arena[r] = static_cast<uint8_t *>(args->arena[r]);
    d.rank = args->rank;
    d.world = args->world;
    d.flag_slot = args->flag_slot;
    d.flag_base = args->flag_base;
    d.data_offset = args->data_offset;
    d.src = args->src;
    d.dst = args->dst;
    d.count = args->count;
    d.src_in_arena = args->src_in_arena;
    d.abort_flag = args->abort_flag;
    d.status = args->status;
    d.timeout_cycles = args->timeout_cycles ? args->timeout_cycles : ~0ull;
    bool is_avg = false;
    const int kop = kernel_op(args->op, is_avg);
    if (kop < 0) return cudaErrorInvalidValue;
    d.avg_divisor = is_avg ? args->world : 0;
    if (pcclKernelP2PStagingBytes(args->count, args->dtype, args->quant_algorithm) > args->data_capacity) return cudaErrorInvalidValue;

    const size_t esz = pcclExtDataTypeSize(static_cast<pcclDataType_t>(args->dtype));
    if (esz == 0) return cudaErrorInvalidValue;
    const uint64_t bytes = args->count * esz;
    const uint64_t n_pack = (bytes + 15) / 16;
    const uint64_t tiles = (n_pack + kThreads - 1) / kThreads;
    int max_blocks = args->max_blocks > 0 ? args->max_blocks : 144;
    if (max_blocks > PCCL_K_MAX_BLOCKS) max_blocks = PCCL_K_MAX_BLOCKS;

    const bool quantised = args->quant_algorithm == QUANT_BLOCK_MINMAX || args->quant_algorithm == QUANT_BLOCK_FP8;
    if (quantised && args->dtype != pcclFloat) return cudaErrorInvalidValue;
    if (!quantised && args->quant_algorithm != QUANT_NONE) return cudaErrorInvalidValue;

    int algorithm = args->algorithm;
    if (algorithm == 0) algorithm = (bytes <= (256u << 10) && !quantised) ? 1 : 2;
    if (quantised) algorithm = 2;
    if (algorithm_out) *algorithm_out = algorithm;

    bool launched = false;
    if (quantised) {
        // per-chunk tiles (a tile = 512 threads x 16 values): work is spread over the chunk of one rank in phase 1
        const uint64_t q_tiles = (args->count + uint64_t(kThreads) * 16 - 1) / (uint64_t(kThreads) * 16);
        const uint64_t chunk_tiles = (q_tiles + args->world - 1) / args->world;
        const int grid = static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(max_blocks, chunk_tiles)));
        dispatch_op(kop, [&](auto op_tag) {
            constexpr int OP = decltype(op_tag)::value;
            if (args->quant_algorithm == QUANT_BLOCK_MINMAX)
                allreduce_two_shot_quant_kernel<QUANT_BLOCK_MINMAX, OP><<<grid, kThreads, 0, stream>>>(d);
            else
                allreduce_two_shot_quant_kernel<QUANT_BLOCK_FP8, OP><<<grid, kThreads, 0, stream>>>(d);
            launched = true;
        });
    } else {
        dispatch_dtype(args->dtype, [&](auto type_tag) {
            using T = decltype(type_tag);
            dispatch_op(kop, [&](auto op_tag) {
                constexpr int OP = decltype(op_tag)::value;
                if (algorithm == 1) {
                    const int grid = static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(max_blocks, tiles)));
                    allreduce_one_shot_kernel<T, OP><<<grid, kThreads, 0, stream>>>(d);
                } else {
                    const uint64_t chunk_tiles = (tiles + args->world - 1) / args->world;
                    const int grid = static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(max_blocks, chunk_tiles)));
                    allreduce_two_shot_kernel<T, OP><<<grid, kThreads, 0, stream>>>(d);
                }
                launched = true;
            });
        });
    }
    if (!launched) return cudaErrorInvalidValue;
    return cudaGetLastError();
}

extern "C" int pcclKernelOuterStepP2P(const pcclKernelP2PArgs_t *args, const pcclKernelOuterStep_t *step, void *stream_v) {
    if (!args || args->world < 1 || args->world > PCCL_K_MAX_RANKS || args->rank >= args->world || args->count == 0) return cudaErrorInvalidValue;
    if (!step || !step->params || !step->inner_in || !step->momentum) return cudaErrorInvalidValue;
    if (args->flag_slot >= static_cast<uint32_t>(kFlagBanks)) return cudaErrorInvalidValue;
    if (pcclKernelP2PStagingBytes(args->count, pcclFloat, QUANT_NONE) > args->data_capacity) return cudaErrorInvalidValue;
    DevArgs d{};
    for (uint32_t r = 0; r < args->world; ++r) d.arena[r] = static_cast<uint8_t *>(args->arena[r]);
    d.rank = args->rank;
    d.world = args->world;
    d.flag_slot = args->flag_slot;
    d.flag_base = args->flag_base;
    d.data_offset = args->data_offset;
    d.count = args->count;
    d.avg_divisor = args->world;
    d.abort_flag = args->abort_flag;
    d.status = args->status;
    d.timeout_cycles = args->timeout_cycles ? args->timeout_cycles : ~0ull;
    OuterStepDev s{};
    s.inner_in = step->inner_in;
    s.params = step->params;
    s.momentum = step->momentum;
    s.inner_out = step->inner_out;
    s.lr = step->lr;
    s.mu = step->mu;
    s.first = step->first;
    s.aligned = ((reinterpret_cast<uintptr_t>(s.inner_in) | reinterpret_cast<uintptr_t>(s.params) | reinterpret_cast<uintptr_t>(s.momentum) |
                  reinterpret_cast<uintptr_t>(s.inner_out)) & 15) == 0;
    const uint64_t n_pack = (args->count + 3) / 4;
    const uint64_t tiles = (n_pack + kThreads - 1) / kThreads;
    int max_blocks = args->max_blocks > 0 ? args->max_blocks : 144;
    if (max_blocks > PCCL_K_MAX_BLOCKS) max_blocks = PCCL_K_MAX_BLOCKS;
    const uint64_t chunk_tiles = (tiles + args->world - 1) / args->world;
    const int grid = static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(max_blocks, chunk_tiles)));
    outer_step_two_shot_kernel<<<grid, kThreads, 0, static_cast<cudaStream_t>(stream_v)>>>(d, s);
    return cudaGetLastError();
}
