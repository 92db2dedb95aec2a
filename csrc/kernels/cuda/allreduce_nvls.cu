// All-reduce through NVSwitch multicast ("NVLS"): the switch performs the reduction
// (multimem.ld_reduce pulls one packet from every GPU's arena and returns the reduced value) and
// the broadcast (multimem.st writes one packet into every GPU's arena).  Per GPU and direction
// about N(1 + 1/P) bytes cross NVLink instead of 2N(P-1)/P for the peer-to-peer two-shot.
//
// No counterpart in the reference (host TCP ring only, ccoip/src/cpp/reduce.cpp:528-784).
//
// Same arena, flag and work-partition scheme as allreduce_p2p.cu.  `mc` is the multicast mapping
// of ALL ranks' arenas (same offsets).  Phases:
//   0  copy the input into the own arena            (skipped for symmetric memory)
//   barrier
//   1  rank r, for the packets of chunk r: ld_reduce over all arenas, AVG scaling, multimem.st
//      of the result back into every arena (in place: only rank r ever touches chunk r here)
//   barrier
//   2  copy the own arena into the output            (skipped for symmetric memory, in place)
// No trailing barrier is needed: a peer can only touch this arena again after the first barrier
// of the NEXT operation, which this rank enters after its copy-out.
#include <cmath>

#include "bulk_copy.cuh"
#include "collective.cuh"
#include "hash_mset.cuh"
#include "outer_step.cuh"
#include "pccl_kernels.h"

namespace pccl::cuda {

constexpr int kNvlsThreads = kThreads;

enum : int { NV_F32 = 0, NV_BF16 = 1, NV_F16 = 2 };

template <int TYPE, int OP> __device__ __forceinline__ uint4 mc_ld_reduce(const void *p) {
    uint4 v;
    if constexpr (TYPE == NV_F32) {
        static_assert(OP == OP_SUM, "multimem.ld_reduce on f32 only adds");
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    } else if constexpr (TYPE == NV_BF16) {
        if constexpr (OP == OP_SUM)
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                         : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
        else if constexpr (OP == OP_MAX)
            asm volatile("multimem.ld_reduce.relaxed.sys.global.max.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                         : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
        else
            asm volatile("multimem.ld_reduce.relaxed.sys.global.min.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                         : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    } else {
        if constexpr (OP == OP_SUM)
            asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                         : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
        else if constexpr (OP == OP_MAX)
            asm volatile("multimem.ld_reduce.relaxed.sys.global.max.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                         : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
        else
            asm volatile("multimem.ld_reduce.relaxed.sys.global.min.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                         : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    }
    return v;
}

__device__ __forceinline__ void mc_store(void *p, uint4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// AVG finalisation: a true division, like the peer-to-peer kernels and the host kernels (x * (1/world) differs in the
// last bit for world sizes that are not powers of two)
template <int TYPE> __device__ __forceinline__ uint4 scale_packet(uint4 v, float div) {
    if constexpr (TYPE == NV_F32) {
        v.x = __float_as_uint(__uint_as_float(v.x) / div);
        v.y = __float_as_uint(__uint_as_float(v.y) / div);
        v.z = __float_as_uint(__uint_as_float(v.z) / div);
        v.w = __float_as_uint(__uint_as_float(v.w) / div);
    } else if constexpr (TYPE == NV_BF16) {
        uint32_t *w = &v.x;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __nv_bfloat162 h = *reinterpret_cast<__nv_bfloat162 *>(&w[i]);
            float2 f = __bfloat1622float2(h);
            h = __floats2bfloat162_rn(f.x / div, f.y / div);
            w[i] = *reinterpret_cast<uint32_t *>(&h);
        }
    } else {
        uint32_t *w = &v.x;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            __half2 h = *reinterpret_cast<__half2 *>(&w[i]);
            float2 f = __half22float2(h);
            h = __floats2half2_rn(f.x / div, f.y / div);
            w[i] = *reinterpret_cast<uint32_t *>(&h);
        }
    }
    return v;
}

// Byte-granular slow paths for unaligned buffers and the last partial packet.  Not inlined on
// purpose: their byte arrays must not drag the callers' packet registers into local memory.
__device__ __noinline__ uint4 load_partial(const uint8_t *src, uint64_t lo, uint64_t hi) {
    alignas(16) uint8_t bytes[16] = {0};
    for (uint64_t b = lo; b < hi; ++b) bytes[b - lo] = src[b];
    return *reinterpret_cast<uint4 *>(bytes);
}
__device__ __noinline__ void store_partial(uint8_t *dst, uint64_t lo, uint64_t hi, uint4 val) {
    alignas(16) uint8_t bytes[16];
    *reinterpret_cast<uint4 *>(bytes) = val;
    for (uint64_t b = lo; b < hi; ++b) dst[b] = bytes[b - lo];
}

// -----------------------------------------------------------------------------------------
// plain NVLS all-reduce.  Input region (data_offset) and result region (red_offset) may differ: symmetric out-of-place
// operands keep their input, and so can be retried after a failure.
// -----------------------------------------------------------------------------------------
template <int TYPE, int OP, int U> struct NvlsAllReduce {
    static __device__ void run(const DevArgs &a) {
        constexpr int ESZ = TYPE == NV_F32 ? 4 : 2;
        constexpr int N = 16 / ESZ;
        __shared__ PeerTable tbl;
        load_peer_table(tbl, a);
        const uint64_t n_pack = (a.count + N - 1) / N;
        const uint64_t per = (n_pack + a.world - 1) / a.world;
        const uint8_t *src = static_cast<const uint8_t *>(a.src);
        uint8_t *dst = static_cast<uint8_t *>(a.dst);
        uint4 *mine = reinterpret_cast<uint4 *>(tbl.base[a.rank]);
        const uint4 *mine_red = reinterpret_cast<const uint4 *>(tbl.red[a.rank]);
        const uint4 *mc_in = reinterpret_cast<const uint4 *>(a.mc + a.data_offset);
        uint4 *mc_out = reinterpret_cast<uint4 *>(a.mc + a.red_offset);
        const uint64_t full_packets = a.count / N; // packets completely inside the user buffers
        const bool aligned = ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0;
        extern __shared__ __align__(128) uint8_t dyn_smem[];
        BulkRing ring;
        if (a.bulk) bulk_ring_init(ring, dyn_smem, 0);
        stamp(a, 0);

        if (!a.src_in_arena && a.bulk) {
            bulk_copy_owned_tiles(ring, reinterpret_cast<uint8_t *>(mine), src, 0, n_pack, 0);
        } else if (!a.src_in_arena) {
            for_owned_packets<4>(0, n_pack, [&](const uint64_t(&v)[4], const bool(&ok)[4]) {
                uint4 tmp[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (!ok[u]) continue;
                    if (aligned && v[u] < full_packets) {
                        tmp[u] = ld_stream_16(src + v[u] * 16);
                    } else {
                        tmp[u] = load_partial(src, v[u] * 16, min(v[u] * 16 + 16, uint64_t(a.count * ESZ)));
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (ok[u]) mine[v[u]] = tmp[u];
            });
        }
        stamp(a, 1);
        if (!peer_barrier<true>(a, tbl, a.flag_base + 1)) {
            if (a.commit_before_write) grid_commit(a, tbl.arena[a.rank], false);
            return;
        }
        stamp(a, 2);

        MsetAcc digest;
        const uint64_t lo = min(uint64_t(a.rank) * per, n_pack), hi = min(lo + per, n_pack);
        const float div = static_cast<float>(a.avg_divisor);
        if (a.dynamic) {
            // Large symmetric operands: SMs differ in how fast they get data through the switch (die, distance to the
            // NVLink ports), and with a static split the whole kernel waits for the slowest block.  Here every warp takes
            // units of 32 x U packets from a counter until the chunk is done; a barrier over all blocks of all ranks follows.
            const int lane = threadIdx.x & 31;
            constexpr uint64_t kUnit = 32ull * U;
            const uint64_t n_units = (hi - lo + kUnit - 1) / kUnit;
            for (;;) {
                unsigned long long unit = 0;
                if (lane == 0) unit = claim_units(a, tbl.arena[a.rank], 1);
                unit = __shfl_sync(0xffffffffu, unit, 0);
                if (unit >= n_units) break;
                uint4 val[U];
                uint64_t v[U];
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    v[u] = lo + unit * kUnit + uint64_t(u) * 32 + lane;
                    if (v[u] < hi) val[u] = mc_ld_reduce<TYPE, OP>(mc_in + v[u]);
                }
#pragma unroll
                for (int u = 0; u < U; ++u) {
                    if (v[u] >= hi) continue;
                    if (a.avg_divisor) val[u] = scale_packet<TYPE>(val[u], div);
                    mc_store(mc_out + v[u], val[u]);
                }
            }
            stamp(a, 3);
            if (!rank_barrier(a, tbl, a.flag_base + 2)) return;
            stamp(a, 4);
            stamp(a, 5);
            return; // dynamic scheduling is only used for symmetric operands: nothing to copy out
        }
        for_owned_packets<U>(lo, hi, [&](const uint64_t(&v)[U], const bool(&ok)[U]) {
            uint4 val[U];
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (ok[u]) val[u] = mc_ld_reduce<TYPE, OP>(mc_in + v[u]);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (!ok[u]) continue;
                if (a.avg_divisor) val[u] = scale_packet<TYPE>(val[u], div);
                mc_store(mc_out + v[u], val[u]);
                if (a.hash_out) digest.add(v[u], val[u], a.count * ESZ);
            }
        });
        stamp(a, 3);
        if (a.hash_out) digest_publish(a, tbl, digest.block_total());
        if (a.commit_before_write) { // in place on an ordinary buffer: all blocks of this rank write the output, or none
            const bool reduced_everywhere = peer_barrier<true>(a, tbl, a.flag_base + 2);
            if (!grid_commit(a, tbl.arena[a.rank], reduced_everywhere)) return;
        } else if (!peer_barrier<true>(a, tbl, a.flag_base + 2)) {
            return;
        }
        if (a.hash_out) digest_collect(a, tbl, a.hash_out);
        stamp(a, 4);

        if (!a.dst_in_arena && a.bulk) {
            // (the bulk engine reads through L2, where the switch delivered the result; nothing stale in L1 to worry about)
            bulk_copy_owned_tiles(ring, dst, reinterpret_cast<const uint8_t *>(mine_red), 0, n_pack, 0);
        } else if (!a.dst_in_arena) {
            for_owned_packets<4>(0, n_pack, [&](const uint64_t(&v)[4], const bool(&ok)[4]) {
                uint4 tmp[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (ok[u]) tmp[u] = ld_peer_16(mine_red + v[u]); // written through the switch: do not trust L1
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (!ok[u]) continue;
                    if (aligned && v[u] < full_packets) {
                        st_stream_16(dst + v[u] * 16, tmp[u]);
                    } else {
                        store_partial(dst, v[u] * 16, min(v[u] * 16 + 16, uint64_t(a.count * ESZ)), tmp[u]);
                    }
                }
            });
        }
        stamp(a, 5);
        // No trailing barrier for staged operands: a peer can only touch this arena's staging area again after the first
        // barrier of the NEXT operation on this flag bank, which this rank enters after its copy-out.  Symmetric operands
        // belong to the caller, who may rewrite them as soon as this kernel ends -- but only the owner of a chunk reads
        // the inputs of that chunk (through the switch), and it is done before it raises the second barrier.
    }
};

// -----------------------------------------------------------------------------------------
// DiLoCo outer step in one launch (fp32, AVG): the staging pass computes the pseudo-gradient, the switch averages
// it, the write-out pass is the Nesterov update.  Nothing the caller owns is written before EVERY block of this rank
// knows that every rank holds the complete average (second barrier + grid-wide commit).
// -----------------------------------------------------------------------------------------
struct NvlsOuter {
    static __device__ void run(const DevArgs &a, const OuterStepDev &s) {
        __shared__ PeerTable tbl;
        load_peer_table(tbl, a);
        const uint64_t n_pack = (a.count + 3) / 4;
        const uint64_t per = (n_pack + a.world - 1) / a.world;
        uint4 *mine = reinterpret_cast<uint4 *>(tbl.base[a.rank]);
        uint4 *mc = reinterpret_cast<uint4 *>(a.mc + a.data_offset);
        stamp(a, 0);

        for_owned_packets<4>(0, n_pack, [&](const uint64_t(&v)[4], const bool(&ok)[4]) {
            uint4 d[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u]) d[u] = outer_delta_packet(s, v[u], a.count);
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u]) mine[v[u]] = d[u];
        });
        stamp(a, 1);
        if (!peer_barrier<true>(a, tbl, a.flag_base + 1)) {
            grid_commit(a, tbl.arena[a.rank], false);
            return;
        }
        stamp(a, 2);

        {
            const uint64_t lo = min(uint64_t(a.rank) * per, n_pack), hi = min(lo + per, n_pack);
            const float div = static_cast<float>(a.world);
            for_owned_packets<4>(lo, hi, [&](const uint64_t(&v)[4], const bool(&ok)[4]) {
                uint4 val[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (ok[u]) val[u] = mc_ld_reduce<NV_F32, OP_SUM>(mc + v[u]);
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (ok[u]) mc_store(mc + v[u], scale_packet<NV_F32>(val[u], div));
            });
        }
        stamp(a, 3);
        const bool reduced_everywhere = peer_barrier<true>(a, tbl, a.flag_base + 2);
        if (!grid_commit(a, tbl.arena[a.rank], reduced_everywhere)) return;
        stamp(a, 4);

        // 4 packets x 3 loads in flight per thread: with one 512-thread block per SM this phase needs ~45 KB
        // outstanding per SM to keep HBM busy
        MsetAcc dig_p, dig_m;
        for_owned_packets<4>(0, n_pack, [&](const uint64_t(&v)[4], const bool(&ok)[4]) {
            uint4 g[4];
            OuterLoaded l[4];
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u]) {
                    g[u] = ld_peer_16(mine + v[u]); // written through the switch: do not trust L1
                    l[u] = outer_load_packet(s, v[u], a.count);
                }
#pragma unroll
            for (int u = 0; u < 4; ++u)
                if (ok[u]) outer_apply_packet(s, v[u], a.count, g[u], l[u], a.hash_out ? &dig_p : nullptr, a.hash_out ? &dig_m : nullptr);
        });
        if (a.hash_out) {
            dig_p.flush(a.hash_out);
            dig_m.flush(a.hash_out + 1);
        }
        stamp(a, 5);
    }
};

// -----------------------------------------------------------------------------------------
// Software-pipelined variant for operands that are NOT in symmetric memory: the three phases of the staged
// kernel above (copy-in / reduce through the switch / copy-out) run CONCURRENTLY on different segments.
// The block is split into two warp groups: the "memory" group (HBM-bound: stages segment i into the arena and
// writes segment i-2 out to the user buffer) and the "network" group (NVLink-bound: reduces segment i-1 through
// the switch).  One cross-GPU barrier per segment keeps the three in step:
//
//   iteration i :   memory group  A(seg i)  [+ C(seg i-2)]   ||   network group  B(seg i-1)      ; barrier
//
// A(seg j) happens-before B(seg j) on every rank (barrier of iteration j), B(seg j) before C(seg j) (barrier of
// iteration j+1), and the three touch disjoint arena regions within one iteration.
// OUTER = true is the DiLoCo outer step: A computes the pseudo-gradient; the Nesterov update must not start before
// the WHOLE average is known to be complete (a failed step may not leave a half-updated model behind), so there
// C is not pipelined: it runs with all warps after the last barrier and the grid-wide commit.
// -----------------------------------------------------------------------------------------
constexpr int kGroupThreads = kNvlsThreads / 2;

template <int U, typename Fn> __device__ __forceinline__ void nvls_group_packets(uint64_t lo, uint64_t hi, int gtid, Fn fn) {
    if (lo >= hi) return;
    constexpr int SUB = kNvlsThreads / kGroupThreads; // a tile of 512 packets takes a 256-thread group two passes
    const uint64_t first_tile = lo / kNvlsThreads;
    const uint64_t t0 = first_tile + ((blockIdx.x + gridDim.x - first_tile % gridDim.x) % gridDim.x);
    for (uint64_t q = 0;; q += U) {
        if ((t0 + (q / SUB) * gridDim.x) * kNvlsThreads >= hi) break;
        uint64_t v[U];
        bool ok[U];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const uint64_t qq = q + u;
            v[u] = (t0 + (qq / SUB) * gridDim.x) * kNvlsThreads + (qq % SUB) * kGroupThreads + gtid;
            ok[u] = v[u] >= lo && v[u] < hi;
        }
        fn(v, ok);
    }
}

struct PipelineGeometry {
    uint32_t n_seg;
    uint64_t seg_pack;
};

template <int TYPE, int OP, bool OUTER> struct NvlsPipelined {
    static __device__ void run(const DevArgs &a, const OuterStepDev &s, const PipelineGeometry &geo) {
        constexpr int ESZ = TYPE == NV_F32 ? 4 : 2;
        constexpr int N = 16 / ESZ;
        __shared__ PeerTable tbl;
        load_peer_table(tbl, a);
        const uint32_t n_seg = geo.n_seg;
        const uint64_t seg_pack = geo.seg_pack;
        const bool memory_group = threadIdx.x < kGroupThreads;
        const int gtid = threadIdx.x % kGroupThreads;
        const uint64_t n_pack = (a.count + N - 1) / N;
        const uint8_t *src = static_cast<const uint8_t *>(a.src);
        uint8_t *dst = static_cast<uint8_t *>(a.dst);
        uint4 *mine = reinterpret_cast<uint4 *>(tbl.base[a.rank]);
        uint4 *mc = reinterpret_cast<uint4 *>(a.mc + a.data_offset);
        const uint64_t full_packets = a.count / N;
        const bool aligned = OUTER || ((reinterpret_cast<uintptr_t>(src) | reinterpret_cast<uintptr_t>(dst)) & 15) == 0;
        const float div = static_cast<float>(a.avg_divisor);
        const uint64_t per = seg_pack / a.world; // the launcher makes segments a multiple of world * tile
        const uint32_t last_iter = OUTER ? n_seg : n_seg + 1;
        MsetAcc digest;
        extern __shared__ __align__(128) uint8_t dyn_smem[];
        BulkRing ring;
        const bool bulk = a.bulk && !OUTER && !a.hash_out;
        if (bulk) bulk_ring_init(ring, dyn_smem, 0);
        stamp(a, 0);

        for (uint32_t i = 0; i <= last_iter; ++i) {
            if (memory_group && bulk) {
                // thread 0 of the memory group drives the copy engine: segment i in, segment i-2 out
                if (i < n_seg) {
                    const uint64_t lo = min(uint64_t(i) * seg_pack, n_pack), hi = min(lo + seg_pack, n_pack);
                    bulk_copy_owned_tiles(ring, reinterpret_cast<uint8_t *>(mine), src, lo, hi, 0);
                }
                if (i >= 2) {
                    const uint64_t lo = min(uint64_t(i - 2) * seg_pack, n_pack), hi = min(lo + seg_pack, n_pack);
                    bulk_copy_owned_tiles(ring, dst, reinterpret_cast<const uint8_t *>(mine), lo, hi, 0);
                }
            } else if (memory_group) {
                if (i < n_seg) { // A: stage segment i
                    const uint64_t lo = min(uint64_t(i) * seg_pack, n_pack), hi = min(lo + seg_pack, n_pack);
                    nvls_group_packets<4>(lo, hi, gtid, [&](const uint64_t(&v)[4], const bool(&ok)[4]) {
                        uint4 tmp[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            if (!ok[u]) continue;
                            if constexpr (OUTER) {
                                tmp[u] = outer_delta_packet(s, v[u], a.count);
                            } else {
                                if (aligned && v[u] < full_packets)
                                    tmp[u] = ld_stream_16(src + v[u] * 16);
                                else
                                    tmp[u] = load_partial(src, v[u] * 16, min(v[u] * 16 + 16, uint64_t(a.count * ESZ)));
                            }
                        }
#pragma unroll
                        for (int u = 0; u < 4; ++u)
                            if (ok[u]) mine[v[u]] = tmp[u];
                    });
                }
                if (!OUTER && i >= 2) { // C: write segment i-2 out
                    const uint64_t lo = min(uint64_t(i - 2) * seg_pack, n_pack), hi = min(lo + seg_pack, n_pack);
                    nvls_group_packets<4>(lo, hi, gtid, [&](const uint64_t(&v)[4], const bool(&ok)[4]) {
                        uint4 tmp[4];
#pragma unroll
                        for (int u = 0; u < 4; ++u)
                            if (ok[u]) tmp[u] = ld_peer_16(mine + v[u]);
#pragma unroll
                        for (int u = 0; u < 4; ++u) {
                            if (!ok[u]) continue;
                            if (aligned && v[u] < full_packets)
                                st_stream_16(dst + v[u] * 16, tmp[u]);
                            else
                                store_partial(dst, v[u] * 16, min(v[u] * 16 + 16, uint64_t(a.count * ESZ)), tmp[u]);
                            if (!OUTER && a.hash_out) digest.add(v[u], tmp[u], a.count * ESZ);
                        }
                    });
                }
            } else if (i >= 1 && i <= n_seg) { // B: reduce this rank's share of segment i-1 through the switch
                const uint64_t slo = min(uint64_t(i - 1) * seg_pack, n_pack), shi = min(slo + seg_pack, n_pack);
                const uint64_t lo = min(slo + uint64_t(a.rank) * per, shi), hi = min(lo + per, shi);
                nvls_group_packets<8>(lo, hi, gtid, [&](const uint64_t(&v)[8], const bool(&ok)[8]) {
                    uint4 val[8];
#pragma unroll
                    for (int u = 0; u < 8; ++u)
                        if (ok[u]) val[u] = mc_ld_reduce<TYPE, OP>(mc + v[u]);
#pragma unroll
                    for (int u = 0; u < 8; ++u) {
                        if (!ok[u]) continue;
                        if (a.avg_divisor) val[u] = scale_packet<TYPE>(val[u], div);
                        mc_store(mc + v[u], val[u]);
                    }
                });
            }
            if (OUTER || i < last_iter) {
                const bool passed = peer_barrier<true>(a, tbl, a.flag_base + 1 + i);
                if constexpr (OUTER) {
                    if (i == last_iter) {
                        if (!grid_commit(a, tbl.arena[a.rank], passed)) return;
                    } else if (!passed) {
                        grid_commit(a, tbl.arena[a.rank], false);
                        return;
                    }
                } else if (!passed) {
                    return;
                }
            }
        }
        if constexpr (OUTER) {
            MsetAcc dig_p, dig_m;
            for_owned_packets<4>(0, n_pack, [&](const uint64_t(&v)[4], const bool(&ok)[4]) {
                uint4 g[4];
                OuterLoaded l[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (ok[u]) {
                        g[u] = ld_peer_16(mine + v[u]);
                        l[u] = outer_load_packet(s, v[u], a.count);
                    }
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (ok[u]) outer_apply_packet(s, v[u], a.count, g[u], l[u], a.hash_out ? &dig_p : nullptr, a.hash_out ? &dig_m : nullptr);
            });
            if (a.hash_out) {
                dig_p.flush(a.hash_out);
                dig_m.flush(a.hash_out + 1);
            }
        } else if (a.hash_out) {
            digest.flush(a.hash_out); // the write-out pass saw every packet of the result
        }
        stamp(a, 5);
    }
};

} // namespace pccl::cuda

using namespace pccl::cuda;

extern "C" int pcclKernelNvlsSupported(int dtype, int op) {
    if (dtype == pcclFloat) return op == pcclSum || op == pcclAvg;
    if (dtype == pcclBFloat16 || dtype == pcclFloat16) return op == pcclSum || op == pcclAvg || op == pcclMax || op == pcclMin;
    return 0;
}

namespace {
bool fill_nvls_args(const pcclKernelNvlsArgs_t *args, int dtype, int rop, DevArgs &d) {
    if (!args || !args->mc_arena) return false;
    const pcclKernelP2PArgs_t &p = args->p2p;
    if (p.world < 2 || p.count == 0 || !pcclKernelNvlsSupported(dtype, rop)) return false;
    if (!fill_dev_args(p, d)) return false;
    d.mc = static_cast<uint8_t *>(args->mc_arena);
    d.avg_divisor = rop == pcclAvg ? p.world : 0;
    const size_t esz = dtype == pcclFloat ? 4 : 2;
    if ((p.count * esz + 15) / 16 * 16 + 256 > p.data_capacity) return false;
    return true;
}
int grid_for(const pcclKernelP2PArgs_t &p, uint64_t chunk_tiles) {
    int max_blocks = p.max_blocks > 0 ? p.max_blocks : 144;
    if (max_blocks > PCCL_K_MAX_BLOCKS) max_blocks = PCCL_K_MAX_BLOCKS;
    return static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(max_blocks, chunk_tiles)));
}
} // namespace

extern "C" int pcclKernelAllReduceNvls(const pcclKernelNvlsArgs_t *args, void *stream_v) {
    if (!args) return cudaErrorInvalidValue;
    const pcclKernelP2PArgs_t &p = args->p2p;
    DevArgs d{};
    if (!fill_nvls_args(args, p.dtype, p.op, d)) return cudaErrorInvalidValue;
    const size_t esz = p.dtype == pcclFloat ? 4 : 2;
    const uint64_t n_pack = (p.count * esz + 15) / 16;
    const uint64_t tiles = (n_pack + kNvlsThreads - 1) / kNvlsThreads;
    const int grid = grid_for(p, (tiles + p.world - 1) / p.world);
    // dynamic work distribution pays off once every block has many units to balance with; symmetric operands only (a staged
    // input would need a barrier over all blocks before the reduction as well); bit 9 of `algorithm` switches it on (measured before it becomes the default)
    d.dynamic = d.src_in_arena && d.dst_in_arena && !d.hash_out && (p.algorithm & 0x200) != 0 &&
                (tiles + p.world - 1) / p.world >= uint64_t(grid) * 8;
    const auto stream = static_cast<cudaStream_t>(stream_v);
    if (d.dynamic) {
        const cudaError_t rc = cudaMemsetAsync(d.arena[d.rank] + claim_word_offset(d.flag_slot), 0, sizeof(unsigned long long), stream);
        if (rc != cudaSuccess) return rc;
    }
    const int op = p.op == pcclMax ? OP_MAX : (p.op == pcclMin ? OP_MIN : OP_SUM);
    // p.algorithm selects a tuning variant: packets kept in flight per thread (default 4; 2 -> 8; 5 -> 16)
    auto launch = [&](auto type_tag, auto op_tag) {
        constexpr int TYPE = decltype(type_tag)::value;
        constexpr int OPV = decltype(op_tag)::value;
        auto go = [&](auto body) {
            using Body = decltype(body);
            const size_t smem = bulk_smem_for(run_rank<Body>, d);
            if (smem == 0) d.bulk = 0;
            run_rank<Body><<<grid, kNvlsThreads, smem, stream>>>(d);
        };
        switch (p.algorithm & 0xff) {
        case 2: go(NvlsAllReduce<TYPE, OPV, 8>{}); break;
        case 5: go(NvlsAllReduce<TYPE, OPV, 16>{}); break;
        default: go(NvlsAllReduce<TYPE, OPV, 4>{}); break;
        }
    };
    using std::integral_constant;
    if (p.dtype == pcclFloat) {
        launch(integral_constant<int, NV_F32>{}, integral_constant<int, OP_SUM>{});
    } else if (p.dtype == pcclBFloat16) {
        if (op == OP_SUM) launch(integral_constant<int, NV_BF16>{}, integral_constant<int, OP_SUM>{});
        else if (op == OP_MAX) launch(integral_constant<int, NV_BF16>{}, integral_constant<int, OP_MAX>{});
        else launch(integral_constant<int, NV_BF16>{}, integral_constant<int, OP_MIN>{});
    } else {
        if (op == OP_SUM) launch(integral_constant<int, NV_F16>{}, integral_constant<int, OP_SUM>{});
        else if (op == OP_MAX) launch(integral_constant<int, NV_F16>{}, integral_constant<int, OP_MAX>{});
        else launch(integral_constant<int, NV_F16>{}, integral_constant<int, OP_MIN>{});
    }
    return cudaGetLastError();
}

extern "C" int pcclKernelOuterStepNvls(const pcclKernelNvlsArgs_t *args, const pcclKernelOuterStep_t *step, void *stream_v) {
    if (!args || !step || !step->params || !step->inner_in || !step->momentum) return cudaErrorInvalidValue;
    const pcclKernelP2PArgs_t &p = args->p2p;
    DevArgs d{};
    if (!fill_nvls_args(args, pcclFloat, pcclAvg, d)) return cudaErrorInvalidValue;
    d.red_offset = d.data_offset;
    d.src_in_arena = d.dst_in_arena = 0;
    const OuterStepDev s = make_outer_step_dev(*step);
    const uint64_t n_pack = (p.count + 3) / 4;
    const uint64_t tiles = (n_pack + kNvlsThreads - 1) / kNvlsThreads;
    const int grid = grid_for(p, (tiles + p.world - 1) / p.world);
    d.bulk = 0;
    run_rank1<NvlsOuter, OuterStepDev><<<grid, kNvlsThreads, 0, static_cast<cudaStream_t>(stream_v)>>>(d, s);
    return cudaGetLastError();
}

// Segments of the pipelined kernel.  With S segments the kernel takes about  T_reduce + (T_in + T_out)/S + (S+2)*b
// (b = one cross-GPU barrier, measured ~15 us; T_in + T_out ~ 0.4 ms for 600 MB), minimal at S = sqrt((T_in+T_out)/b):
// 5 for 600 MB, growing with the square root of the size (measured on 4xB200: 3:1.54  4:1.52  5:1.51  6:1.53  8:1.58
// 12:1.71 ms against 1.73 ms unpipelined, profiles/tune4_kernel_pipeline.jsonl).  Segments are a whole number of
// (world x tile) packets.
static void pipeline_geometry(uint64_t n_pack, uint32_t world, int requested, uint32_t &n_seg, uint64_t &seg_pack) {
    uint64_t want = requested > 0 ? static_cast<uint64_t>(requested)
                                  : static_cast<uint64_t>(5.2 * std::sqrt(static_cast<double>(n_pack) * 16.0 / 6.0e8) + 0.5);
    want = std::min<uint64_t>(24, std::max<uint64_t>(2, want)); // segments + 2 barrier values must fit the caller's flag budget (32)
    const uint64_t unit = uint64_t(world) * kNvlsThreads;
    seg_pack = ((n_pack + want - 1) / want + unit - 1) / unit * unit;
    n_seg = static_cast<uint32_t>((n_pack + seg_pack - 1) / seg_pack);
}

extern "C" int pcclKernelAllReduceNvlsPipelined(const pcclKernelNvlsArgs_t *args, const pcclKernelOuterStep_t *step, int segments,
                                                void *stream_v) {
    if (!args) return cudaErrorInvalidValue;
    const pcclKernelP2PArgs_t &p = args->p2p;
    const int dtype = step ? static_cast<int>(pcclFloat) : p.dtype;
    const int rop = step ? static_cast<int>(pcclAvg) : p.op;
    if (step && (!step->params || !step->inner_in || !step->momentum)) return cudaErrorInvalidValue;
    DevArgs d{};
    if (!fill_nvls_args(args, dtype, rop, d)) return cudaErrorInvalidValue;
    d.red_offset = d.data_offset; // staged: the segments are reduced in place in the arena
    d.src_in_arena = d.dst_in_arena = 0;
    const size_t esz = dtype == pcclFloat ? 4 : 2;
    OuterStepDev s{};
    if (step) s = make_outer_step_dev(*step);
    const uint64_t n_pack = (p.count * esz + 15) / 16;
    PipelineGeometry geo{};
    pipeline_geometry(n_pack, p.world, segments, geo.n_seg, geo.seg_pack);
    // the network group's share of one segment decides how many blocks have work in every iteration
    const int grid = grid_for(p, std::max<uint64_t>(1, geo.seg_pack / p.world / kNvlsThreads));
    const auto stream = static_cast<cudaStream_t>(stream_v);
    const int op = rop == pcclMax ? OP_MAX : (rop == pcclMin ? OP_MIN : OP_SUM);
    auto go = [&](auto body) {
        using Body = decltype(body);
        const size_t smem = bulk_smem_for(run_rank2<Body, OuterStepDev, PipelineGeometry>, d);
        if (smem == 0) d.bulk = 0;
        run_rank2<Body, OuterStepDev, PipelineGeometry><<<grid, kNvlsThreads, smem, stream>>>(d, s, geo);
    };
    if (step) {
        go(NvlsPipelined<NV_F32, OP_SUM, true>{});
    } else if (dtype == pcclFloat) {
        go(NvlsPipelined<NV_F32, OP_SUM, false>{});
    } else if (dtype == pcclBFloat16) {
        if (op == OP_SUM) go(NvlsPipelined<NV_BF16, OP_SUM, false>{});
        else if (op == OP_MAX) go(NvlsPipelined<NV_BF16, OP_MAX, false>{});
        else go(NvlsPipelined<NV_BF16, OP_MIN, false>{});
    } else {
        if (op == OP_SUM) go(NvlsPipelined<NV_F16, OP_SUM, false>{});
        else if (op == OP_MAX) go(NvlsPipelined<NV_F16, OP_MAX, false>{});
        else go(NvlsPipelined<NV_F16, OP_MIN, false>{});
    }
    return cudaGetLastError();
}
