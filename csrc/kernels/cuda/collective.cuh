// Shared machinery of the collective kernels (allreduce_p2p.cu, allreduce_nvls.cu): the argument block as the
// device sees it, the control-page layout, the cross-GPU barrier, the per-rank grid-wide commit, the work partition.
//
// Control page (first pcclKernelP2PControlBytes() bytes of every arena):
//   [0, 2 MiB)            kFlagBanks banks of 256-byte lines, one line per block: flags[src_rank] (u64 x 16, cross-GPU
//                         barrier) followed by digests[src_rank] (u64 x 16, partial content hashes exchanged between ranks)
//   [2 MiB, +4 KiB)       per-bank grid-commit words (arrive word, fail word), one 128-byte line per bank
// Flag values only grow (the caller derives them from an operation counter), so values left behind by an aborted
// operation can never satisfy a later wait.
#pragma once
#include <atomic>
#include <cstdlib>

#include "common.cuh"
#include "pccl_kernels.h"

namespace pccl::cuda {

constexpr int kThreads = 512;
constexpr int kFlagBanks = 32;
static_assert(kFlagBanks == PCCL_K_FLAG_BANKS, "header and kernels disagree");
constexpr size_t kFlagLineBytes = 256;
constexpr size_t kFlagBankBytes = size_t(PCCL_K_MAX_BLOCKS) * kFlagLineBytes;
constexpr size_t kGridSyncOffset = size_t(kFlagBanks) * kFlagBankBytes;
constexpr size_t kGridSyncStride = 128;
constexpr size_t kControlBytes = 4u << 20;
static_assert(kGridSyncOffset + kFlagBanks * kGridSyncStride <= kControlBytes, "control page too small");
static_assert(PCCL_K_MAX_RANKS * 8 <= 128, "flag line too small");

struct DevArgs {
    uint8_t *arena[PCCL_K_MAX_RANKS]; // unicast mappings of every rank's arena
    uint8_t *mc;                      // multicast mapping of the arenas (NVLS kernels only)
    uint32_t rank, world, flag_slot;
    uint32_t serial;                  // launch counter of this process, tags the grid-commit words
    uint64_t flag_base;               // barrier values are flag_base + 1..
    uint64_t data_offset;             // where peers read this operation's input in every arena
    uint64_t red_offset;              // where chunk owners leave their reduced chunks in every arena
    const void *src;
    void *dst;
    uint64_t count;
    uint64_t avg_divisor; // 0: none
    int src_in_arena;     // the input already is arena[rank] + data_offset (symmetric memory)
    int dst_in_arena;     // the output IS arena[rank] + red_offset (symmetric memory): no second store
    int trailing_barrier; // close with a barrier (nobody may reuse peer-visible memory a peer still reads)
    int fence_system;     // barriers start with a block-wide system fence (tuning/diagnosis; the release stores suffice)
    int bulk;             // pure copy phases use the bulk asynchronous copy engine (bulk_copy.cuh); needs its dynamic shared memory
    int dynamic;          // NVLS: warps take work units from a shared counter instead of a static tile assignment
    int commit_before_write; // the output aliases the (staged) input: write it only once every block of this rank knows
                             // that the reduction completed everywhere, so a failed operation leaves the input intact
    volatile uint32_t *abort_flag;
    uint32_t *status;
    uint64_t timeout_cycles;
    unsigned long long *stamps; // optional (host-mapped): %globaltimer at the phase boundaries, written by block 0
    // fused content hash of the output (optional, see hash_tile.cuh): digest lands in *hash_out
    unsigned long long *hash_out;
};

__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void stamp(const DevArgs &a, int i) {
    if (a.stamps && blockIdx.x == 0 && threadIdx.x == 0) a.stamps[i] = global_timer_ns();
}

// Peer base pointers live in shared memory: they are indexed with run-time rank numbers, and a run-time index
// into kernel parameters would be demoted to a per-thread stack copy.
struct PeerTable {
    uint8_t *arena[PCCL_K_MAX_RANKS];
    uint8_t *base[PCCL_K_MAX_RANKS]; // arena + data_offset
    uint8_t *red[PCCL_K_MAX_RANKS];  // arena + red_offset
};
__device__ __forceinline__ void load_peer_table(PeerTable &t, const DevArgs &a) {
    if (threadIdx.x < PCCL_K_MAX_RANKS) {
        const bool live = threadIdx.x < a.world;
        t.arena[threadIdx.x] = live ? a.arena[threadIdx.x] : nullptr;
        t.base[threadIdx.x] = live ? a.arena[threadIdx.x] + a.data_offset : nullptr;
        t.red[threadIdx.x] = live ? a.arena[threadIdx.x] + a.red_offset : nullptr;
    }
    __syncthreads();
}
constexpr int kPeerGroup = 8; // peers whose loads are kept in flight together

__device__ __forceinline__ uint64_t *flag_ptr(const DevArgs &a, uint8_t *arena_base, uint32_t block, uint32_t writer) {
    return reinterpret_cast<uint64_t *>(arena_base + a.flag_slot * kFlagBankBytes + size_t(block) * kFlagLineBytes) + writer;
}
__device__ __forceinline__ unsigned long long *digest_ptr(const DevArgs &a, uint8_t *arena_base, uint32_t block, uint32_t writer) {
    return reinterpret_cast<unsigned long long *>(arena_base + a.flag_slot * kFlagBankBytes + size_t(block) * kFlagLineBytes + 128) + writer;
}

__device__ __forceinline__ uint32_t ld_acquire_gpu_u32(const uint32_t *p) {
    uint32_t v;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long ld_acquire_gpu_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// Cross-GPU barrier among the same-numbered blocks of all ranks.  Returns false when the operation was aborted or
// timed out (uniform across the block).  FENCE_SYSTEM: multimem stores of this block must be performed before the
// flag is raised.
template <bool FENCE_SYSTEM = false> __device__ bool peer_barrier(const DevArgs &a, const PeerTable &t, uint64_t value, bool record_failure = true) {
    __shared__ int s_fail;
    if (threadIdx.x == 0) s_fail = 0;
    // Stores of this block (plain or multimem) become visible to a peer that acquires the flag through the release stores
    // below, which are cumulative over everything the block barrier orders before them.  A fence by every thread costs
    // ~10 us per barrier on B200 and is only kept as a diagnostic switch.
    if constexpr (FENCE_SYSTEM) {
        if (a.fence_system) __threadfence_system();
    }
    __syncthreads(); // this block's earlier writes are ordered before the flag stores below
    if (threadIdx.x < a.world) {
        const uint32_t peer = threadIdx.x;
        st_flag_release(flag_ptr(a, t.arena[peer], blockIdx.x, a.rank), value);
        const uint64_t *mine = flag_ptr(a, t.arena[a.rank], blockIdx.x, peer);
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

// Barrier among ALL blocks of ALL ranks (the per-block barrier above only ties same-numbered blocks together): every
// block arrives at a counter in its own rank's control page; the last one raises this rank's flag in line 0 of every
// arena; everybody waits for all ranks' flags there.  Used where work is handed out dynamically, so that any block may
// have produced what any other block consumes next.  words[2] = (serial << 32) | arrivals.
static __device__ bool rank_barrier(const DevArgs &a, const PeerTable &t, uint64_t value) {
    __shared__ int s_last, s_fail;
    __syncthreads();
    if (threadIdx.x == 0) {
        s_fail = 0;
        __threadfence_system(); // one thread: cumulative over the block's stores, multimem included
        auto *word = reinterpret_cast<unsigned long long *>(t.arena[a.rank] + kGridSyncOffset + a.flag_slot * kGridSyncStride) + 2;
        const unsigned long long mine = static_cast<unsigned long long>(a.serial) << 32;
        unsigned long long old = ld_acquire_gpu_u64(word), now;
        for (;;) {
            now = (old >> 32) == a.serial ? old + 1 : (mine | 1ull);
            const unsigned long long seen = atomicCAS(word, old, now);
            if (seen == old) break;
            old = seen;
        }
        s_last = (now & 0xffffffffull) == gridDim.x;
    }
    __syncthreads();
    if (threadIdx.x < a.world) {
        const uint32_t peer = threadIdx.x;
        if (s_last) st_flag_release(flag_ptr(a, t.arena[peer], 0, a.rank), value);
        const uint64_t *mine = flag_ptr(a, t.arena[a.rank], 0, peer);
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
    if (fail != 0 && threadIdx.x == 0) atomicMax(a.status, static_cast<uint32_t>(fail));
    return fail == 0;
}

// Work counter for dynamically scheduled phases (words[3], zeroed on the stream by the launcher before the kernel):
// returns the first of `n` units claimed for the caller.
static __device__ __forceinline__ unsigned long long claim_units(const DevArgs &a, uint8_t *my_arena, uint32_t n) {
    auto *word = reinterpret_cast<unsigned long long *>(my_arena + kGridSyncOffset + a.flag_slot * kGridSyncStride) + 3;
    return atomicAdd(word, static_cast<unsigned long long>(n));
}
inline size_t claim_word_offset(uint32_t flag_slot) { return kGridSyncOffset + size_t(flag_slot) * kGridSyncStride + 24; }

// Per-rank, grid-wide agreement: returns true in EVERY block of this rank's grid iff every block called with
// ok == true; otherwise false everywhere.  Used where a kernel is about to write memory the caller owns and must
// do so either in all blocks or in none (the outer-step update).  All blocks of a collective kernel are co-resident
// by construction (they already spin on their peers), so waiting for siblings cannot deadlock.
// words[0] = (serial << 32) | arrivals, words[1] = serial of a launch in which some block failed.
static __device__ bool grid_commit(const DevArgs &a, uint8_t *my_arena, bool ok) {
    __shared__ int s_commit;
    if (threadIdx.x == 0) {
        auto *words = reinterpret_cast<unsigned long long *>(my_arena + kGridSyncOffset + a.flag_slot * kGridSyncStride);
        const unsigned long long mine = static_cast<unsigned long long>(a.serial) << 32;
        int res = ok ? 1 : 0;
        if (!ok) {
            atomicExch(words + 1, static_cast<unsigned long long>(a.serial));
        } else {
            unsigned long long old = ld_acquire_gpu_u64(words);
            for (;;) {
                const unsigned long long want = (old >> 32) == a.serial ? old + 1 : (mine | 1ull);
                const unsigned long long seen = atomicCAS(words, old, want);
                if (seen == old) break;
                old = seen;
            }
            __threadfence();
            const long long t0 = clock64();
            uint32_t spins = 0;
            for (;;) {
                const unsigned long long w = ld_acquire_gpu_u64(words);
                if ((w >> 32) == a.serial && (w & 0xffffffffull) >= gridDim.x) break;
                if (ld_acquire_gpu_u64(words + 1) == a.serial) {
                    res = 0;
                    break;
                }
                if ((++spins & 0x3ffu) == 0 && (*a.abort_flag != 0 || static_cast<uint64_t>(clock64() - t0) > a.timeout_cycles)) {
                    // re-check the count once more: if everybody arrived meanwhile the decision is "commit" for all
                    const unsigned long long w2 = ld_acquire_gpu_u64(words);
                    if ((w2 >> 32) == a.serial && (w2 & 0xffffffffull) >= gridDim.x) break;
                    atomicExch(words + 1, static_cast<unsigned long long>(a.serial));
                    atomicMax(a.status, *a.abort_flag != 0 ? 1u : 2u);
                    res = 0;
                    break;
                }
            }
        }
        s_commit = res;
    }
    __syncthreads();
    return s_commit != 0;
}

// Content hash of a collective's result without a pass of its own: every rank hashes the chunk it reduces (the values are
// in registers anyway), same-numbered blocks swap their partial sums through the control pages around the barrier that
// follows, and every rank ends up with the sum over all chunks = the mset64 raw sum of the complete result.
//   digest_publish   before the barrier: this block's partial goes to every rank's line
//   digest_collect   after the barrier: sum what the twins left here, add it to *out (one atomic per block)
__device__ __forceinline__ void digest_publish(const DevArgs &a, const PeerTable &t, unsigned long long block_total) {
    if (threadIdx.x < a.world) {
        unsigned long long *p = digest_ptr(a, t.arena[threadIdx.x], blockIdx.x, a.rank);
        asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(block_total) : "memory");
    }
}
__device__ __forceinline__ void digest_collect(const DevArgs &a, const PeerTable &t, unsigned long long *out) {
    if (threadIdx.x < 32) {
        unsigned long long v = 0;
        if (threadIdx.x < a.world) {
            const unsigned long long *p = digest_ptr(a, t.arena[a.rank], blockIdx.x, threadIdx.x);
            asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if (threadIdx.x == 0 && v != 0) atomicAdd(out, v);
    }
}

// A kernel body is a struct with a static run(); the two wrappers below turn it into the real kernel (one rank per
// launch, ranks on different GPUs) and into the emulated form used by single-GPU tests: ONE cooperative launch whose
// grid is (blocks, ranks), every y-slice playing one rank on its own arena.  (Ranks that wait for each other must
// never be separate launches on one GPU -- nothing guarantees that they run at the same time.)
// (fixed arities instead of a parameter pack: the address of a kernel is needed for cudaFuncSetAttribute, and a
// specialisation whose trailing pack could still grow has no unique address)
template <typename Body> __global__ void __launch_bounds__(kThreads, 1) run_rank(const DevArgs a) { Body::run(a); }
template <typename Body, typename E1> __global__ void __launch_bounds__(kThreads, 1) run_rank1(const DevArgs a, const E1 e1) { Body::run(a, e1); }
template <typename Body, typename E1, typename E2> __global__ void __launch_bounds__(kThreads, 1) run_rank2(const DevArgs a, const E1 e1, const E2 e2) {
    Body::run(a, e1, e2);
}
template <typename Body, typename... Extra> __global__ void __launch_bounds__(kThreads, 1) run_emulated(const DevArgs *ranks, const Extra *...extra) {
    Body::run(ranks[blockIdx.y], extra[blockIdx.y]...);
}

// Iterates over the packets of [lo, hi) that belong to this block, U tiles at a time.  Tile t (kThreads packets)
// belongs to block t % gridDim.x ON EVERY RANK AND IN EVERY PHASE, so a block only ever consumes data the
// same-numbered block of a peer produced.  fn(v[], ok[]) gets up to U packet indices.
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

// host side: fills the device argument block from the C ABI structure (common part of all launchers)
inline uint32_t next_launch_serial() {
    static std::atomic<uint32_t> serial{1};
    uint32_t v = serial.fetch_add(1, std::memory_order_relaxed);
    if (v == 0) v = serial.fetch_add(1, std::memory_order_relaxed); // 0 is what a fresh control page holds
    return v;
}
inline bool fill_dev_args(const pcclKernelP2PArgs_t &p, DevArgs &d) {
    if (p.world < 1 || p.world > PCCL_K_MAX_RANKS || p.rank >= p.world) return false;
    if (p.flag_slot >= static_cast<uint32_t>(kFlagBanks)) return false;
    if (p.data_offset % 16 != 0 || p.red_offset % 16 != 0) return false;
    d = DevArgs{};
    for (uint32_t r = 0; r < p.world; ++r) d.arena[r] = static_cast<uint8_t *>(p.arena[r]);
    d.rank = p.rank;
    d.world = p.world;
    d.flag_slot = p.flag_slot;
    d.serial = next_launch_serial();
    d.flag_base = p.flag_base;
    d.data_offset = p.data_offset;
    d.red_offset = p.red_offset ? p.red_offset : p.data_offset;
    d.src = p.src;
    d.dst = p.dst;
    d.count = p.count;
    d.src_in_arena = p.src_in_arena;
    d.dst_in_arena = p.dst_in_arena;
    d.trailing_barrier = p.no_trailing_barrier ? 0 : 1;
    d.fence_system = (p.algorithm & 0x100) ? 1 : 0;
    {
        // bulk copies move whole 16-byte packets between 16-byte aligned addresses; small operands are latency-bound anyway
        static const bool enabled = [] {
            const char *v = std::getenv("PCCL_BULK_COPY");
            return !(v && v[0] == '0');
        }();
        const uint64_t bytes = p.count * pcclExtDataTypeSize(static_cast<pcclDataType_t>(p.dtype));
        const auto misaligned = (reinterpret_cast<uintptr_t>(p.src) | reinterpret_cast<uintptr_t>(p.dst)) & 15;
        d.bulk = enabled && (p.algorithm & 0x400) == 0 && !misaligned && bytes % 16 == 0 && bytes >= (1u << 20) && p.quant_algorithm == 0;
    }
    d.commit_before_write = (!p.src_in_arena && !p.dst_in_arena && p.src == p.dst && p.src != nullptr) ? 1 : 0;
    d.abort_flag = p.abort_flag;
    d.status = p.status;
    d.timeout_cycles = p.timeout_cycles ? p.timeout_cycles : ~0ull;
    d.stamps = reinterpret_cast<unsigned long long *>(p.stamps);
    d.hash_out = reinterpret_cast<unsigned long long *>(p.hash_out);
    return true;
}

} // namespace pccl::cuda
