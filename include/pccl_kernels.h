/*
 * pccl_kernels.h -- C ABI of the sm_100a kernels inside libpccl.so.
 *
 * These entry points are what the communicator itself calls; they are exported so that tests,
 * benchmarks and other runtimes (PyTorch through ctypes, for example) can drive single kernels.
 * Every launcher is asynchronous on the given stream (a cudaStream_t passed as void*) unless
 * stated otherwise, and returns 0 on success or a cudaError_t value.
 *
 * Data type / operator numbering is that of pcclDataType_t / pcclRedOp_t in pccl.h.
 */
#ifndef PCCL_KERNELS_H_
#define PCCL_KERNELS_H_

#include "pccl.h"

#ifdef __cplusplus
extern "C" {
#endif

#define PCCL_K_MAX_RANKS 16      /* peers in one NVLink domain */
#define PCCL_K_MAX_BLOCKS 256    /* thread blocks of one collective kernel */
#define PCCL_K_FLAG_BANKS 32     /* independent barrier-flag banks in the control page (flag_slot < this) */
#define PCCL_K_QUANT_BLOCK 128   /* elements sharing one scale in the block formats */

/* ---------------------------------------------------------------------------------------------
 * Content hashes (reference kernels: ccoip/src/cuda/simplehash_cuda.cu:95-156, no GPU CRC in the
 * reference).  `scratch` must point to pcclKernelHashScratchBytes() bytes of device memory that
 * was zeroed once; the kernels leave it zeroed again.  The result is written to *out_device
 * (device memory, 8 bytes).
 * ------------------------------------------------------------------------------------------- */
PCCL_EXPORT size_t pcclKernelHashScratchBytes(void);
PCCL_EXPORT int pcclKernelSimpleHash(const void *data, size_t n_bytes, void *scratch, uint64_t *out_device, void *stream);
PCCL_EXPORT int pcclKernelCrc32c(const void *data, size_t n_bytes, void *scratch, uint64_t *out_device, void *stream);

/* mset64, the order-independent hash type (no counterpart in the reference; definition in csrc/kernels/mset_formula.hpp):
 * writes the RAW sum to *out_device; pcclExtMsetFinalize(raw, n_bytes) turns a raw sum into the hash value.  The collective
 * kernels produce the same raw sums as a by-product (hash_out of pcclKernelP2PArgs_t). */
PCCL_EXPORT int pcclKernelMset64Raw(const void *data, size_t n_bytes, uint64_t *out_device, void *stream);
PCCL_EXPORT uint64_t pcclExtMsetFinalize(uint64_t raw_sum, uint64_t n_bytes);

/* Convenience: allocates nothing itself -- uses a per-device scratch the library keeps -- runs the
 * hash on an internal stream and returns the value on the host (synchronous). */
PCCL_EXPORT int pcclKernelHashSync(const void *data, size_t n_bytes, int hash_type, uint64_t *out_host); /* 0 simplehash, 1 CRC32C, 2 mset64 */

/* The same for `count` buffers with ONE wait at the end (shared-state synchronisation of a parameter list). */
PCCL_EXPORT int pcclKernelHashManySync(const void *const *data, const size_t *n_bytes, size_t count, int hash_type, uint64_t *out_host);

/* Same hash, for callers that may not allocate device memory (open stream capture, allocator that already
 * owns the whole GPU, paths that cannot afford cudaMalloc's device-wide synchronisation).  Has no
 * counterpart in the reference.  The scratch is a 4 KiB static device array inside the library, shared
 * by all calls on a device: calls whose kernels overlap on the GPU race, so order them.  The first
 * call on each device resolves the array's address and should happen before any stream capture.
 * Inputs larger than pcclKernelSimpleHashNoAllocMaxBytes() (2^32 - 1 sixteen-byte packets) are refused
 * with cudaErrorInvalidValue and a message on stderr.  Asynchronous on `stream`; the 8-byte result
 * lands in *out_device (device memory supplied by the caller). */
PCCL_EXPORT int pcclKernelSimpleHashNoAlloc(const void *data, size_t n_bytes, uint64_t *out_device, void *stream);
PCCL_EXPORT size_t pcclKernelSimpleHashNoAllocMaxBytes(void);

/* ---------------------------------------------------------------------------------------------
 * Element-wise kernels of the TCP data plane on device buffers (reference CPU versions:
 * ccoip/src/cpp/reduce_kernels.cpp:56-215,437-510; quantize_kernels.cpp:24-217).
 * ------------------------------------------------------------------------------------------- */
/* dst[i] = dst[i] (op) src[i] */
PCCL_EXPORT int pcclKernelReduce(void *dst, const void *src, size_t count, int dtype, int op, void *stream);
/* dst[i] /= divisor, in the element type */
PCCL_EXPORT int pcclKernelScaleInv(void *dst, size_t count, int dtype, uint64_t divisor, void *stream);

/* Block formats: out = [meta for ceil(count/128) blocks][count code bytes].
 *   algorithm pcclQuantBlockMinMax: meta = (min,max) fp32 pairs, codes uint8
 *   algorithm pcclQuantBlockFp8:    meta = fp32 scale,           codes e4m3        */
PCCL_EXPORT size_t pcclKernelQuantizedBytes(size_t count, int algorithm);
PCCL_EXPORT int pcclKernelQuantize(const float *src, void *out, size_t count, int algorithm, void *stream);
/* dst[i] = dst[i] (op) dequant(in[i]); op = -1 stores the dequantised value */
PCCL_EXPORT int pcclKernelDequantizeReduce(float *dst, const void *in, size_t count, int algorithm, int op, void *stream);

/* ---------------------------------------------------------------------------------------------
 * Peer-to-peer all-reduce over NVLink (no counterpart in the reference, which reduces on the
 * host: ccoip/src/cpp/reduce.cpp:528-784).  One launch per rank; all ranks pass the same
 * geometry.  arena[r] is rank r's arena as mapped into the calling process.
 * ------------------------------------------------------------------------------------------- */
typedef struct pcclKernelP2PArgs_t {
    void *arena[PCCL_K_MAX_RANKS];
    uint32_t rank, world;
    uint32_t flag_slot;     /* which flag bank of the control page */
    uint64_t flag_base;     /* barrier values are flag_base + 1.. ; must grow from call to call */
    uint64_t data_offset;   /* byte offset of the staging region inside every arena */
    uint64_t data_capacity; /* bytes available there */
    const void *src;        /* local input */
    void *dst;              /* local output */
    uint64_t count;         /* elements */
    int dtype, op;
    int quant_algorithm;    /* pcclQuantNone, pcclQuantBlockMinMax, pcclQuantBlockFp8 */
    int algorithm;          /* 0 = choose, 1 = one-shot, 2 = two-shot */
    int max_blocks;         /* SM budget, 0 = default */
    int src_in_arena;       /* src already is arena[rank]+data_offset (symmetric memory): skip the copy-in */
    volatile uint32_t *abort_flag; /* host-mapped; non-zero makes the kernel give up */
    uint32_t *status;       /* host-mapped; 0 ok, 1 aborted, 2 timed out */
    uint64_t timeout_cycles;
    /* -- optional members; zero selects the classic behaviour -- */
    uint64_t red_offset;     /* where chunk owners leave their reduced chunks in every arena; 0 = data_offset */
    int dst_in_arena;        /* dst IS arena[rank]+red_offset (symmetric memory): results are written once, there */
    int no_trailing_barrier; /* one-shot only: the caller alternates between two staging areas from call to call on this
                                flag bank, so the closing "everybody has read" barrier is not needed */
    uint64_t *stamps;        /* host-mapped, 8 words: %globaltimer (ns) at the phase boundaries, written by block 0 */
    uint64_t *hash_out;      /* device memory, zeroed by the caller: raw mset64 sum of the output (two-shot, NVLS; the outer
                                step writes two words: new parameters, new momentum).  Finalise with pcclExtMsetFinalize. */
} pcclKernelP2PArgs_t;

/* Returns which algorithm ran (1 or 2) in *algorithm_out. */
PCCL_EXPORT int pcclKernelAllReduceP2P(const pcclKernelP2PArgs_t *args, int *algorithm_out, void *stream);
/* Single-GPU emulation for tests: ranks[0..n_ranks) are the argument blocks of n_ranks = world ranks, each with its OWN
 * arena[] table (all arenas in this GPU's memory).  Runs them as ONE cooperative kernel with grid (blocks, n_ranks) --
 * kernels of ranks that wait for each other must not be separate launches on one GPU.  Synchronous; blocks * n_ranks must
 * fit the GPU (keep max_blocks small). */
PCCL_EXPORT int pcclKernelEmulatedAllReduceP2P(const pcclKernelP2PArgs_t *ranks, int n_ranks, int *algorithm_out, void *stream);
/* Bytes of arena staging needed for `count` elements. */
PCCL_EXPORT size_t pcclKernelP2PStagingBytes(uint64_t count, int dtype, int quant_algorithm);
/* Size of the control page at the start of every arena. */
PCCL_EXPORT size_t pcclKernelP2PControlBytes(void);

/* NVSwitch multicast variant: mc_base is the multicast mapping of all arenas.  Only
 * sum/min/max/avg on f32, f16, bf16 (what multimem.ld_reduce offers). */
typedef struct pcclKernelNvlsArgs_t {
    pcclKernelP2PArgs_t p2p;
    void *mc_arena; /* multicast view */
} pcclKernelNvlsArgs_t;
PCCL_EXPORT int pcclKernelAllReduceNvls(const pcclKernelNvlsArgs_t *args, void *stream);
/* Software-pipelined form for operands outside symmetric memory: copy-in of segment i, switch reduction of segment
 * i-1 and copy-out of segment i-2 run concurrently on two warp groups of every block (one cross-GPU barrier per
 * segment).  `step` non-NULL runs the DiLoCo outer step instead of a plain all-reduce (see pcclKernelOuterStepNvls
 * below; its update phase is not pipelined so that a failed step leaves the tensors untouched).
 * `segments` 0 = about 32 MiB each.  flag_base must leave room for segments + 2 barrier values. */
struct pcclKernelOuterStep_t;
PCCL_EXPORT int pcclKernelAllReduceNvlsPipelined(const pcclKernelNvlsArgs_t *args, const struct pcclKernelOuterStep_t *step, int segments,
                                                 void *stream);
/* 1 when multimem.ld_reduce can do (dtype, op). */
PCCL_EXPORT int pcclKernelNvlsSupported(int dtype, int op);

/* ---------------------------------------------------------------------------------------------
 * DiLoCo outer step as ONE kernel per rank: pseudo-gradient, fp32 AVG all-reduce, Nesterov-SGD update
 * (the reference's example does the three on the host with torch ops around pcclAllReduce:
 * python/examples/nanogpt_diloco/sync_diloco.py:480-530).
 *   staging   arena[v] = params[v] - inner_in[v]                (replaces the copy-in)
 *   reduce    as in the plain kernels, result left in the arenas only
 *   update    g = averaged pseudo-gradient; momentum = first ? g : mu*momentum + g;
 *             params -= lr*(g + mu*momentum); inner_out = params (replaces the copy-out)
 * params/momentum/inner_out are only written after the barrier that proves every rank holds the complete
 * average, so an aborted operation leaves them untouched.  Geometry (arena, flags, count, rank, world,
 * abort word, status, timeout, max_blocks) comes from the all-reduce argument block; its src, dst,
 * dtype, op, quant_algorithm, algorithm and src_in_arena members are ignored.
 * ------------------------------------------------------------------------------------------- */
typedef struct pcclKernelOuterStep_t {
    const float *inner_in;
    float *params;
    float *momentum;
    float *inner_out; /* may be NULL */
    float lr, mu;
    int first; /* momentum holds nothing yet */
} pcclKernelOuterStep_t;
/* The two halves as separate element-wise passes (used when the fused kernel cannot run):
 * out = params - inner, and the update above with `grad` in place of g. */
PCCL_EXPORT int pcclKernelPseudoGrad(float *out, const float *params, const float *inner, size_t count, void *stream);
PCCL_EXPORT int pcclKernelOuterNesterov(float *params, float *momentum, const float *grad, float *inner_out, size_t count, float lr, float mu,
                                        int first, void *stream);
PCCL_EXPORT int pcclKernelOuterStepP2P(const pcclKernelP2PArgs_t *args, const pcclKernelOuterStep_t *step, void *stream);
PCCL_EXPORT int pcclKernelOuterStepNvls(const pcclKernelNvlsArgs_t *args, const pcclKernelOuterStep_t *step, void *stream);
/* Single-GPU emulation of pcclKernelOuterStepP2P, see pcclKernelEmulatedAllReduceP2P. */
PCCL_EXPORT int pcclKernelEmulatedOuterStepP2P(const pcclKernelP2PArgs_t *ranks, const pcclKernelOuterStep_t *steps, int n_ranks, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* PCCL_KERNELS_H_ */
