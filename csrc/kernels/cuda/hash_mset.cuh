// Device side of the mset64 content hash (definition: csrc/kernels/mset_formula.hpp): a per-thread accumulator that
// a kernel feeds with every packet it writes, and a block-level flush that adds the block's share to a 64-bit word
// in device memory (the raw sum; the host finalises it).
#pragma once
#include "../mset_formula.hpp"
#include "common.cuh"

namespace pccl::cuda {

struct MsetAcc {
    unsigned long long sum = 0;

    // packet `v` of a buffer of n_bytes bytes; bytes past the end of the buffer count as zero
    __device__ __forceinline__ void add(uint64_t v, uint4 p, uint64_t n_bytes) {
        const uint64_t end = v * 16 + 16;
        if (end > n_bytes) {
            const uint32_t keep = v * 16 >= n_bytes ? 0u : static_cast<uint32_t>(n_bytes - v * 16); // 1..15 bytes
            if (keep == 0) return;
            uint32_t wds[4] = {p.x, p.y, p.z, p.w};
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int lo = 4 * i;
                if (static_cast<int>(keep) <= lo)
                    wds[i] = 0;
                else if (static_cast<int>(keep) < lo + 4)
                    wds[i] &= (1u << (8 * (keep - lo))) - 1u;
            }
            p = make_uint4(wds[0], wds[1], wds[2], wds[3]);
        }
        sum += mset::term(v, p.x, p.y, p.z, p.w);
    }

    // all threads of the block call; returns the block's sum in every thread and clears the accumulator
    __device__ __forceinline__ unsigned long long block_total() {
        __shared__ unsigned long long s_part[33];
        unsigned long long v = sum;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
        __syncthreads();
        if (lane == 0) s_part[warp] = v;
        __syncthreads();
        if (warp == 0) {
            v = lane < (blockDim.x + 31) / 32 ? s_part[lane] : 0ull;
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
            if (lane == 0) s_part[32] = v;
        }
        __syncthreads();
        sum = 0;
        return s_part[32];
    }

    // all threads of the block call; one atomic per block
    __device__ __forceinline__ void flush(unsigned long long *out) {
        const unsigned long long total = block_total();
        if (threadIdx.x == 0 && total != 0) atomicAdd(out, total);
    }
};

} // namespace pccl::cuda
