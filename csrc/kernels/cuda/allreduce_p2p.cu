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
#include <tuple>
#include <type_traits>

#include "bulk_copy.cuh"
#include "collective.cuh"
#include "hash_mset.cuh"
#include "outer_step.cuh"
#include "pccl_kernels.h"
#include "quant.cuh"

namespace pccl::cuda {

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
//   phase 0  copy own input into own arena                              (skipped for symmetric input)
//   barrier
//   phase 1  rank r reduces chunk r: reads chunk r of every arena's input region, reduces in rank order, applies
//            AVG, writes the result to the "reduced" region of its arena (and to its output when that is elsewhere)
//   barrier
//   phase 2  reads the reduced chunks of all other ranks from their arenas into the output
//   barrier  (nobody may overwrite peer-visible memory a peer is still reading)
// -----------------------------------------------------------------------------------------
template <typename T, int OP> struct TwoShot {
    static __device__ void run(const DevArgs &a) {
        constexpr int N = 16 / sizeof(T);
        __shared__ PeerTable tbl;
        load_peer_table(tbl, a);
        const uint64_t n_pack = (a.count + N - 1) / N;
        const uint64_t per = (n_pack + a.world - 1) / a.world; // packets per chunk
        const T *src = static_cast<const T *>(a.src);
        T *dst = static_cast<T *>(a.dst);
        const bool src_aligned = (reinterpret_cast<uintptr_t>(src) & 15) == 0, dst_aligned = (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
        uint4 *mine = reinterpret_cast<uint4 *>(tbl.base[a.rank]);
        uint4 *mine_red = reinterpret_cast<uint4 *>(tbl.red[a.rank]);
        MsetAcc digest;
        extern __shared__ __align__(128) uint8_t dyn_smem[];
        BulkRing ring;
        if (a.bulk) bulk_ring_init(ring, dyn_smem, 0);
        stamp(a, 0);

        // phase 0: stage the input where the peers can read it
        if (!a.src_in_arena && a.bulk) {
            bulk_copy_owned_tiles(ring, reinterpret_cast<uint8_t *>(mine), static_cast<const uint8_t *>(a.src), 0, n_pack, 0);
        } else if (!a.src_in_arena) {
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
        stamp(a, 1);
        if (!peer_barrier(a, tbl, a.flag_base + 1)) {
            if (a.commit_before_write) grid_commit(a, tbl.arena[a.rank], false);
            return;
        }
        stamp(a, 2);

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
                    mine_red[v[u]] = out;
                    if (!a.dst_in_arena && !a.commit_before_write) store_output(dst, v[u], a.count, dst_aligned, out);
                    if (a.hash_out) digest.add(v[u], out, a.count * sizeof(T));
                }
            });
        }
        stamp(a, 3);
        if (a.hash_out) digest_publish(a, tbl, digest.block_total());
        if (a.commit_before_write) {
            // in place on an ordinary buffer: nothing of the caller's has been written so far (the own chunk waits in the
            // arena too); from here on the result is complete on every rank, and all blocks of this rank write or none
            const bool reduced_everywhere = peer_barrier(a, tbl, a.flag_base + 2);
            if (!grid_commit(a, tbl.arena[a.rank], reduced_everywhere)) return;
        } else if (!peer_barrier(a, tbl, a.flag_base + 2)) {
            return;
        }
        if (a.hash_out) digest_collect(a, tbl, a.hash_out);
        stamp(a, 4);

        // phase 2: collect the other ranks' reduced chunks, starting with the right-hand neighbour so that at any
        // moment the ranks pull from different peers
        for (uint32_t k = a.commit_before_write ? 0 : 1; k < a.world; ++k) {
            const uint32_t peer = (a.rank + k) % a.world;
            const uint4 *from = reinterpret_cast<const uint4 *>(tbl.red[peer]);
            const uint64_t lo = min(uint64_t(peer) * per, n_pack), hi = min(lo + per, n_pack);
            for_owned_packets<4>(lo, hi, [&](const uint64_t(&v)[4], const bool(&ok)[4]) {
                uint4 tmp[4];
#pragma unroll
                for (int u = 0; u < 4; ++u)
                    if (ok[u]) tmp[u] = ld_peer_16(from + v[u]);
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (!ok[u]) continue;
                    if (a.dst_in_arena)
                        mine_red[v[u]] = tmp[u];
                    else
                        store_output(dst, v[u], a.count, dst_aligned, tmp[u]);
                }
            });
        }
        stamp(a, 5);
        peer_barrier(a, tbl, a.flag_base + 3);
        stamp(a, 6);
    }
};

// -----------------------------------------------------------------------------------------
// one-shot, unquantised (latency regime): phase 0, barrier, every rank reduces everything in a fixed order
// (bit-identical everywhere).  What follows depends on who owns the memory the peers read:
//   * staged input, double-buffered by the caller (trailing_barrier == 0): nothing -- the NEXT operation on this
//     flag bank uses the other half of the staging area, and nobody can start the one after that before every peer
//     has passed that operation's first barrier, i.e. has finished reading this one;
//   * staged input, single-buffered: a closing barrier;
//   * symmetric input: the caller may rewrite its buffer as soon as this kernel ends, so a barrier "everybody has
//     read everything" comes first; when the output IS the input (in place) the results wait in registers for it.
// -----------------------------------------------------------------------------------------
constexpr int kOneShotRegs = 4; // packets per thread the in-place form can hold back
template <typename T, int OP> struct OneShot {
    static __device__ void run(const DevArgs &a) {
        constexpr int N = 16 / sizeof(T);
        __shared__ PeerTable tbl;
        load_peer_table(tbl, a);
        const uint64_t n_pack = (a.count + N - 1) / N;
        const T *src = static_cast<const T *>(a.src);
        T *dst = static_cast<T *>(a.dst);
        const bool src_aligned = (reinterpret_cast<uintptr_t>(src) & 15) == 0, dst_aligned = (reinterpret_cast<uintptr_t>(dst) & 15) == 0;
        uint4 *mine = reinterpret_cast<uint4 *>(tbl.base[a.rank]);
        stamp(a, 0);
        if (!a.src_in_arena) {
            for_owned_packets<1>(0, n_pack, [&](const uint64_t(&v)[1], const bool(&ok)[1]) {
                if (ok[0]) mine[v[0]] = load_input(src, v[0], a.count, src_aligned);
            });
        }
        if (!peer_barrier(a, tbl, a.flag_base + 1)) {
            if (a.commit_before_write) grid_commit(a, tbl.arena[a.rank], false);
            return;
        }
        stamp(a, 2);
        auto reduce_packet = [&](uint64_t v) -> uint4 {
            typename Elem<T>::acc_t acc[N];
            for (uint32_t r0 = 0; r0 < a.world; r0 += kPeerGroup) { // fixed order: bit-identical on every rank
                uint4 raw[kPeerGroup];
#pragma unroll
                for (int g = 0; g < kPeerGroup; ++g)
                    if (r0 + g < a.world) raw[g] = ld_peer_16(reinterpret_cast<const uint4 *>(tbl.base[r0 + g]) + v);
#pragma unroll
                for (int g = 0; g < kPeerGroup; ++g)
                    if (r0 + g < a.world) accumulate<T, OP>(acc, raw[g], r0 + g == 0);
            }
            return finish<T>(acc, a.avg_divisor);
        };
        const bool in_place = a.dst_in_arena && a.red_offset == a.data_offset;
        if (in_place) {
            // the launcher guarantees that one pass of kOneShotRegs tiles per block covers the operand
            uint4 held[kOneShotRegs];
            uint64_t hv[kOneShotRegs];
            bool hok[kOneShotRegs];
#pragma unroll
            for (int u = 0; u < kOneShotRegs; ++u) {
                hv[u] = (uint64_t(blockIdx.x) + uint64_t(u) * gridDim.x) * kThreads + threadIdx.x;
                hok[u] = hv[u] < n_pack;
                if (hok[u]) held[u] = reduce_packet(hv[u]);
            }
            if (!peer_barrier(a, tbl, a.flag_base + 2)) return; // every rank has read every input
#pragma unroll
            for (int u = 0; u < kOneShotRegs; ++u)
                if (hok[u]) mine[hv[u]] = held[u];
            stamp(a, 6);
            return;
        }
        // in place on an ordinary buffer: every block of this rank must have seen all inputs before any of them writes
        if (a.commit_before_write && !grid_commit(a, tbl.arena[a.rank], true)) return;
        uint4 *out_sym = reinterpret_cast<uint4 *>(tbl.red[a.rank]);
        for_owned_packets<1>(0, n_pack, [&](const uint64_t(&v)[1], const bool(&ok)[1]) {
            if (!ok[0]) return;
            const uint4 r = reduce_packet(v[0]);
            if (a.dst_in_arena)
                out_sym[v[0]] = r;
            else
                store_output(dst, v[0], a.count, dst_aligned, r);
        });
        if (a.trailing_barrier) peer_barrier(a, tbl, a.flag_base + 2);
        stamp(a, 6);
    }
};

// -----------------------------------------------------------------------------------------
// two-shot with block-scaled wire format (fp32 in memory, 1 byte per element on NVLink)
//   staging layout: [u32 codes, one per float4][block records, one per 32 float4]
// -----------------------------------------------------------------------------------------
template <int ALGO, int OP> struct TwoShotQuant {
  static __device__ void run(const DevArgs &a) {
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
    if (!peer_barrier(a, tbl, a.flag_base + 1)) {
        if (a.commit_before_write) grid_commit(a, tbl.arena[a.rank], false);
        return;
    }

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
            if (!a.commit_before_write) {
                float y[16]; // the owner keeps exactly what the others will decode
                dequantize16<ALGO>(packed, lo, hi, y);
                store16(v[0], y);
            }
        });
    }
    if (a.commit_before_write) { // in place: write nothing of the caller's before the reduction is complete everywhere
        const bool reduced_everywhere = peer_barrier(a, tbl, a.flag_base + 2);
        if (!grid_commit(a, tbl.arena[a.rank], reduced_everywhere)) return;
    } else if (!peer_barrier(a, tbl, a.flag_base + 2)) {
        return;
    }

    // phase 2: fetch the other chunks (in place: the own one too) as codes, decode into the output
    for (uint32_t k = a.commit_before_write ? 0 : 1; k < a.world; ++k) {
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
};

// -----------------------------------------------------------------------------------------
// DiLoCo outer step in one launch (fp32, AVG), peer-to-peer flavour of outer_step_nvls_kernel:
//   phase 0  arena = params - inner
//   phase 1  owner of a chunk averages it over all arenas, result goes to its arena ONLY
//   phase 2  every rank walks all chunks (its own included) and applies the Nesterov update
// The caller's tensors are first written in phase 2, when every rank provably holds the whole
// average; a failure of the closing barrier (a peer died after that point) is therefore not an
// error of this operation and is not recorded.
// -----------------------------------------------------------------------------------------
struct OuterTwoShot {
    static __device__ void run(const DevArgs &a, const OuterStepDev &s) {
        __shared__ PeerTable tbl;
        load_peer_table(tbl, a);
        const uint64_t n_pack = (a.count + 3) / 4;
        const uint64_t per = (n_pack + a.world - 1) / a.world;
        uint4 *mine = reinterpret_cast<uint4 *>(tbl.base[a.rank]);
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
        if (!peer_barrier(a, tbl, a.flag_base + 1)) {
            grid_commit(a, tbl.arena[a.rank], false);
            return;
        }
        stamp(a, 2);

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
        stamp(a, 3);
        // The update below writes the caller's tensors.  It must happen in every block of this rank or in none: a block
        // that passed the barrier while a sibling gave up (abort raised in between) would leave params half-updated.
        const bool reduced_everywhere = peer_barrier(a, tbl, a.flag_base + 2);
        if (!grid_commit(a, tbl.arena[a.rank], reduced_everywhere)) return;
        stamp(a, 4);

        MsetAcc dig_p, dig_m;
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
                    if (ok[u]) outer_apply_packet(s, v[u], a.count, g[u], l[u], a.hash_out ? &dig_p : nullptr, a.hash_out ? &dig_m : nullptr);
            });
        }
        stamp(a, 5);
        if (a.hash_out) {
            dig_p.flush(a.hash_out);
            dig_m.flush(a.hash_out + 1);
        }
        peer_barrier(a, tbl, a.flag_base + 3, false);
        stamp(a, 6);
    }
};

} // namespace pccl::cuda

using namespace pccl::cuda;

extern "C" size_t pcclKernelP2PControlBytes(void) { return kControlBytes; }

extern "C" size_t pcclKernelP2PStagingBytes(uint64_t count, int dtype, int quant_algorithm) {
    if (quant_algorithm == QUANT_BLOCK_MINMAX || quant_algorithm == QUANT_BLOCK_FP8 || quant_algorithm == QUANT_BLOCK_ZPS) {
        const uint64_t n_blocks = (count + PCCL_K_QUANT_BLOCK - 1) / PCCL_K_QUANT_BLOCK;
        const uint64_t mb = quant_algorithm == QUANT_BLOCK_FP8 ? 4 : 8;
        return n_blocks * PCCL_K_QUANT_BLOCK + n_blocks * mb + 256;
    }
    const size_t esz = pcclExtDataTypeSize(static_cast<pcclDataType_t>(dtype));
    return (count * esz + 15) / 16 * 16 + 256;
}

namespace {

// Launches Body either as the real kernel (n_ranks == 1 entry, this process' rank) or -- tests on a single GPU -- as ONE
// cooperative kernel whose grid is (grid, n_ranks).  The emulated form is synchronous.
template <typename Body, typename... Extra>
int launch_body(const DevArgs *ranks, int n_ranks, bool emulated, int grid, cudaStream_t stream, const Extra *...extra) {
    if (!emulated) {
        DevArgs d = ranks[0];
        if constexpr (sizeof...(Extra) == 0) {
            const size_t smem = bulk_smem_for(run_rank<Body>, d);
            if (smem == 0) d.bulk = 0;
            run_rank<Body><<<grid, kThreads, smem, stream>>>(d);
        } else {
            const size_t smem = bulk_smem_for(run_rank1<Body, Extra...>, d);
            if (smem == 0) d.bulk = 0;
            run_rank1<Body, Extra...><<<grid, kThreads, smem, stream>>>(d, extra[0]...);
        }
        return cudaGetLastError();
    }
    DevArgs *d_ranks = nullptr;
    cudaError_t rc = cudaMalloc(&d_ranks, sizeof(DevArgs) * n_ranks);
    if (rc != cudaSuccess) return rc;
    std::tuple<Extra *...> d_extra{};
    auto upload = [&](auto *&dev, const auto *host) {
        using E = std::remove_cv_t<std::remove_pointer_t<decltype(host)>>;
        if (rc != cudaSuccess) return;
        rc = cudaMalloc(&dev, sizeof(E) * n_ranks);
        if (rc == cudaSuccess) rc = cudaMemcpyAsync(dev, host, sizeof(E) * n_ranks, cudaMemcpyHostToDevice, stream);
    };
    DevArgs no_bulk[PCCL_K_MAX_RANKS]; // the emulated (test) launch keeps the register copy loops
    for (int i = 0; i < n_ranks; ++i) {
        no_bulk[i] = ranks[i];
        no_bulk[i].bulk = 0;
    }
    rc = cudaMemcpyAsync(d_ranks, no_bulk, sizeof(DevArgs) * n_ranks, cudaMemcpyHostToDevice, stream);
    std::apply([&](auto *&...dev) { (upload(dev, extra), ...); }, d_extra);
    if (rc == cudaSuccess) {
        std::apply(
            [&](auto *...dev) {
                const DevArgs *ranks_arg = d_ranks;
                void *params[] = {&ranks_arg, (void *) &dev...};
                rc = cudaLaunchCooperativeKernel(reinterpret_cast<void *>(run_emulated<Body, Extra...>), dim3(grid, n_ranks), dim3(kThreads), params, 0,
                                                 stream);
            },
            d_extra);
    }
    const cudaError_t sync = cudaStreamSynchronize(stream);
    if (rc == cudaSuccess) rc = sync;
    cudaFree(d_ranks);
    std::apply([&](auto *...dev) { (cudaFree(dev), ...); }, d_extra);
    return rc;
}

int launch_allreduce(const pcclKernelP2PArgs_t *ranks, int n_ranks, bool emulated, int *algorithm_out, cudaStream_t stream) {
    if (!ranks || n_ranks < 1 || n_ranks > PCCL_K_MAX_RANKS) return cudaErrorInvalidValue;
    const pcclKernelP2PArgs_t *args = &ranks[0];
    bool is_avg = false;
    const int kop = kernel_op(args->op, is_avg);
    if (kop < 0) return cudaErrorInvalidValue;
    const size_t esz = pcclExtDataTypeSize(static_cast<pcclDataType_t>(args->dtype));
    if (esz == 0) return cudaErrorInvalidValue;
    DevArgs d[PCCL_K_MAX_RANKS];
    for (int i = 0; i < n_ranks; ++i) {
        const pcclKernelP2PArgs_t &p = ranks[i];
        if (!fill_dev_args(p, d[i])) return cudaErrorInvalidValue;
        if (p.count != args->count || p.dtype != args->dtype || p.op != args->op || p.quant_algorithm != args->quant_algorithm ||
            p.world != args->world)
            return cudaErrorInvalidValue;
        d[i].avg_divisor = is_avg ? p.world : 0;
        if (pcclKernelP2PStagingBytes(p.count, p.dtype, p.quant_algorithm) > p.data_capacity) return cudaErrorInvalidValue;
        if (emulated) d[i].serial = d[0].serial; // every emulated rank owns its arena; one serial keeps the logs readable
    }
    const uint64_t bytes = args->count * esz;
    const uint64_t n_pack = (bytes + 15) / 16;
    const uint64_t tiles = (n_pack + kThreads - 1) / kThreads;
    int max_blocks = args->max_blocks > 0 ? args->max_blocks : 144;
    if (max_blocks > PCCL_K_MAX_BLOCKS) max_blocks = PCCL_K_MAX_BLOCKS;

    const bool quantised = args->quant_algorithm == QUANT_BLOCK_MINMAX || args->quant_algorithm == QUANT_BLOCK_FP8 ||
                           args->quant_algorithm == QUANT_BLOCK_ZPS;
    if (quantised && args->dtype != pcclFloat) return cudaErrorInvalidValue;
    if (!quantised && args->quant_algorithm != QUANT_NONE) return cudaErrorInvalidValue;

    int algorithm = args->algorithm;
    if (algorithm == 0) algorithm = (bytes <= (256u << 10) && !quantised) ? 1 : 2;
    if (quantised) algorithm = 2;
    int grid = 1;
    if (algorithm == 1) {
        grid = static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(max_blocks, tiles)));
        // in place on symmetric memory the results wait in registers for the "everybody has read" barrier: one pass only
        const bool in_place_symmetric = d[0].dst_in_arena && d[0].red_offset == d[0].data_offset;
        if (in_place_symmetric && tiles > uint64_t(grid) * kOneShotRegs) algorithm = 2;
        if (d[0].hash_out) algorithm = 2; // the fused digest is produced by the two-shot write-out
    }
    if (algorithm_out) *algorithm_out = algorithm;

    int rc = cudaErrorInvalidValue;
    if (quantised) {
        // per-chunk tiles (a tile = 512 threads x 16 values): work is spread over the chunk of one rank in phase 1
        const uint64_t q_tiles = (args->count + uint64_t(kThreads) * 16 - 1) / (uint64_t(kThreads) * 16);
        const uint64_t chunk_tiles = (q_tiles + args->world - 1) / args->world;
        grid = static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(max_blocks, chunk_tiles)));
        dispatch_op(kop, [&](auto op_tag) {
            constexpr int OP = decltype(op_tag)::value;
            if (args->quant_algorithm == QUANT_BLOCK_MINMAX)
                rc = launch_body<TwoShotQuant<QUANT_BLOCK_MINMAX, OP>>(d, n_ranks, emulated, grid, stream);
            else if (args->quant_algorithm == QUANT_BLOCK_ZPS)
                rc = launch_body<TwoShotQuant<QUANT_BLOCK_ZPS, OP>>(d, n_ranks, emulated, grid, stream);
            else
                rc = launch_body<TwoShotQuant<QUANT_BLOCK_FP8, OP>>(d, n_ranks, emulated, grid, stream);
        });
    } else {
        if (algorithm != 1) {
            const uint64_t chunk_tiles = (tiles + args->world - 1) / args->world;
            grid = static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(max_blocks, chunk_tiles)));
        }
        dispatch_dtype(args->dtype, [&](auto type_tag) {
            using T = decltype(type_tag);
            dispatch_op(kop, [&](auto op_tag) {
                constexpr int OP = decltype(op_tag)::value;
                if (algorithm == 1)
                    rc = launch_body<OneShot<T, OP>>(d, n_ranks, emulated, grid, stream);
                else
                    rc = launch_body<TwoShot<T, OP>>(d, n_ranks, emulated, grid, stream);
            });
        });
    }
    return rc;
}

int launch_outer_step(const pcclKernelP2PArgs_t *ranks, const pcclKernelOuterStep_t *steps, int n_ranks, bool emulated, cudaStream_t stream) {
    if (!ranks || !steps || n_ranks < 1 || n_ranks > PCCL_K_MAX_RANKS) return cudaErrorInvalidValue;
    DevArgs d[PCCL_K_MAX_RANKS];
    OuterStepDev s[PCCL_K_MAX_RANKS];
    for (int i = 0; i < n_ranks; ++i) {
        const pcclKernelP2PArgs_t &p = ranks[i];
        const pcclKernelOuterStep_t &step = steps[i];
        if (p.count == 0 || p.count != ranks[0].count || !step.params || !step.inner_in || !step.momentum) return cudaErrorInvalidValue;
        if (pcclKernelP2PStagingBytes(p.count, pcclFloat, QUANT_NONE) > p.data_capacity) return cudaErrorInvalidValue;
        if (!fill_dev_args(p, d[i])) return cudaErrorInvalidValue;
        d[i].avg_divisor = p.world;
        d[i].red_offset = d[i].data_offset;
        d[i].src_in_arena = d[i].dst_in_arena = 0;
        if (emulated) d[i].serial = d[0].serial;
        d[i].bulk = 0; // the outer step's staging and update passes do arithmetic; no pure copy in it
        s[i] = make_outer_step_dev(step);
    }
    const uint64_t n_pack = (ranks[0].count + 3) / 4;
    const uint64_t tiles = (n_pack + kThreads - 1) / kThreads;
    int max_blocks = ranks[0].max_blocks > 0 ? ranks[0].max_blocks : 144;
    if (max_blocks > PCCL_K_MAX_BLOCKS) max_blocks = PCCL_K_MAX_BLOCKS;
    const uint64_t chunk_tiles = (tiles + ranks[0].world - 1) / ranks[0].world;
    const int grid = static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(max_blocks, chunk_tiles)));
    return launch_body<OuterTwoShot, OuterStepDev>(d, n_ranks, emulated, grid, stream, s);
}

} // namespace

extern "C" int pcclKernelAllReduceP2P(const pcclKernelP2PArgs_t *args, int *algorithm_out, void *stream_v) {
    return launch_allreduce(args, 1, false, algorithm_out, static_cast<cudaStream_t>(stream_v));
}

extern "C" int pcclKernelEmulatedAllReduceP2P(const pcclKernelP2PArgs_t *ranks, int n_ranks, int *algorithm_out, void *stream_v) {
    if (!ranks || n_ranks < 1 || static_cast<uint32_t>(n_ranks) != ranks[0].world) return cudaErrorInvalidValue;
    return launch_allreduce(ranks, n_ranks, true, algorithm_out, static_cast<cudaStream_t>(stream_v));
}

extern "C" int pcclKernelOuterStepP2P(const pcclKernelP2PArgs_t *args, const pcclKernelOuterStep_t *step, void *stream_v) {
    return launch_outer_step(args, step, 1, false, static_cast<cudaStream_t>(stream_v));
}

extern "C" int pcclKernelEmulatedOuterStepP2P(const pcclKernelP2PArgs_t *ranks, const pcclKernelOuterStep_t *steps, int n_ranks, void *stream_v) {
    if (!ranks || n_ranks < 1 || static_cast<uint32_t>(n_ranks) != ranks[0].world) return cudaErrorInvalidValue;
    return launch_outer_step(ranks, steps, n_ranks, true, static_cast<cudaStream_t>(stream_v));
}
