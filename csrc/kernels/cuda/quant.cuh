// Block-scaled quantisation on the device: 128 consecutive fp32 values share one scale record.
// One warp handles one block: every lane owns 4 consecutive values (one float4 in, one packed
// u32 of code bytes out), the block statistics come from warp shuffles.
// Numerics are identical to the host oracle in csrc/kernels/cpu/quantize.cpp (block_minmax_* and
// block_fp8_*), operation by operation.
#pragma once
#include "common.cuh"

namespace pccl::cuda {

// pcclQuantizationAlgorithm_t values; QUANT_BLOCK_ZPS is how pcclQuantZeroPointScale (2) runs on the device plane: uint8 codes, one
// (scale, zero point) record per 128 elements -- scale = (max-min)/255, zp = round(-min/scale), q = clamp(round(x/scale) + zp),
// x' = (q - zp) * scale; a constant block is stored as scale = value, q = 1, so it survives exactly (the reference's anchor:
// 42 + 43 must come out as exactly 85, ccoip/tests/end_to_end/test_all_reduce.cpp:326-399).  The reference takes these
// formulas from the piquant library per ring chunk (ccoip/src/cpp/quantize.cpp:22-57); here they are per block.
enum : int { QUANT_NONE = 0, QUANT_BLOCK_ZPS = 2, QUANT_BLOCK_MINMAX = 3, QUANT_BLOCK_FP8 = 4 };

template <int ALGO> struct QuantMeta;
template <> struct QuantMeta<QUANT_BLOCK_MINMAX> {
    static constexpr int kBytes = 8;
};
template <> struct QuantMeta<QUANT_BLOCK_FP8> {
    static constexpr int kBytes = 4;
};
template <> struct QuantMeta<QUANT_BLOCK_ZPS> {
    static constexpr int kBytes = 8;
};

__device__ __forceinline__ float warp_min(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

__device__ __forceinline__ uint32_t fp8_encode(float x) {
    return static_cast<uint32_t>(__nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3));
}
__device__ __forceinline__ float fp8_decode(uint32_t c) {
    const __half_raw h = __nv_cvt_fp8_to_halfraw(static_cast<__nv_fp8_storage_t>(c & 0xffu), __NV_E4M3);
    return __half2float(*reinterpret_cast<const __half *>(&h));
}

// Quantises the 4 values of this lane; all 32 lanes of the warp must call together.  `valid`
// tells how many of the 4 values exist (tail of the buffer); missing ones do not influence the
// statistics.  Returns the packed codes, writes the block record through `meta` (lane 0 only).
template <int ALGO> __device__ __forceinline__ uint32_t quantize4(const float (&x)[4], int valid, void *meta, int lane) {
    if constexpr (ALGO == QUANT_BLOCK_MINMAX) {
        float mn = 3.402823466e+38f, mx = -3.402823466e+38f;
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (i < valid) {
                mn = fminf(mn, x[i]);
                mx = fmaxf(mx, x[i]);
            }
        mn = warp_min(mn);
        mx = warp_max(mx);
        if (lane == 0) *reinterpret_cast<float2 *>(meta) = make_float2(mn, mx);
        const float inv = mx > mn ? 255.0f / (mx - mn) : 0.0f;
        uint32_t packed = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float r = rintf((x[i] - mn) * inv);
            const uint32_t c = i < valid ? static_cast<uint32_t>(fminf(255.0f, fmaxf(0.0f, r))) : 0u;
            packed |= c << (8 * i);
        }
        return packed;
    } else {
        float amax = 0.0f;
#pragma unroll
        for (int i = 0; i < 4; ++i)
            if (i < valid) amax = fmaxf(amax, fabsf(x[i]));
        amax = warp_max(amax);
        const float scale = amax > 0.0f ? amax / 448.0f : 1.0f;
        if (lane == 0) *reinterpret_cast<float *>(meta) = scale;
        const float inv = 1.0f / scale;
        uint32_t packed = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) packed |= (i < valid ? fp8_encode(x[i] * inv) : 0u) << (8 * i);
        return packed;
    }
}

// meta_lo / meta_hi: the raw 32-bit words of the block record (hi unused for fp8)
template <int ALGO> __device__ __forceinline__ void dequantize4(uint32_t packed, uint32_t meta_lo, uint32_t meta_hi, float (&out)[4]) {
    if constexpr (ALGO == QUANT_BLOCK_MINMAX) {
        const float mn = __uint_as_float(meta_lo), mx = __uint_as_float(meta_hi);
        const float step = (mx - mn) * (1.0f / 255.0f);
#pragma unroll
        for (int i = 0; i < 4; ++i) out[i] = fmaf(static_cast<float>((packed >> (8 * i)) & 0xffu), step, mn);
    } else if constexpr (ALGO == QUANT_BLOCK_ZPS) {
        const float scale = __uint_as_float(meta_lo), zp = static_cast<float>(static_cast<int>(meta_hi));
#pragma unroll
        for (int i = 0; i < 4; ++i) out[i] = (static_cast<float>((packed >> (8 * i)) & 0xffu) - zp) * scale;
    } else {
        const float scale = __uint_as_float(meta_lo);
#pragma unroll
        for (int i = 0; i < 4; ++i) out[i] = fp8_decode(packed >> (8 * i)) * scale;
    }
}


// ---- 16 elements per thread, 8 lanes per block -------------------------------------------------
// Used by the fused all-reduce: one 16-byte code packet per thread keeps the peer loads wide, and
// the block statistics need 3 shuffle steps per 16 elements instead of 5 per 4.
// Only the 8 lanes of the caller's group take part (other groups of the warp may have left
// because their packets lie beyond the end of the buffer), so the mask names exactly that group.
__device__ __forceinline__ uint32_t group8_mask() { return 0xffu << ((threadIdx.x & 31u) & ~7u); }
__device__ __forceinline__ float group8_min(float v) {
    const uint32_t mask = group8_mask();
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) v = fminf(v, __shfl_xor_sync(mask, v, o));
    return v;
}
__device__ __forceinline__ float group8_max(float v) {
    const uint32_t mask = group8_mask();
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(mask, v, o));
    return v;
}

// `valid` = how many of the 16 values exist.  All 8 lanes of a group must call together.
// Returns the block record in (rec_lo, rec_hi) on every lane of the group.
template <int ALGO>
__device__ __forceinline__ uint4 quantize16(const float (&x)[16], int valid, uint32_t &rec_lo, uint32_t &rec_hi) {
    uint32_t w[4] = {0, 0, 0, 0};
    if constexpr (ALGO == QUANT_BLOCK_MINMAX) {
        float mn = 3.402823466e+38f, mx = -3.402823466e+38f;
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if (i < valid) {
                mn = fminf(mn, x[i]);
                mx = fmaxf(mx, x[i]);
            }
        mn = group8_min(mn);
        mx = group8_max(mx);
        rec_lo = __float_as_uint(mn), rec_hi = __float_as_uint(mx);
        const float inv = mx > mn ? 255.0f / (mx - mn) : 0.0f;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const float r = rintf((x[i] - mn) * inv);
            const uint32_t c = i < valid ? static_cast<uint32_t>(fminf(255.0f, fmaxf(0.0f, r))) : 0u;
            w[i >> 2] |= c << (8 * (i & 3));
        }
    } else if constexpr (ALGO == QUANT_BLOCK_ZPS) {
        float mn = 3.402823466e+38f, mx = -3.402823466e+38f;
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if (i < valid) {
                mn = fminf(mn, x[i]);
                mx = fmaxf(mx, x[i]);
            }
        mn = group8_min(mn);
        mx = group8_max(mx);
        float scale;
        int zp;
        if (!(mx > mn)) { // constant block: code 1 times scale = the value (code 0 for zeros)
            scale = mn != 0.0f ? mn : 1.0f;
            zp = 0;
        } else {
            scale = (mx - mn) * (1.0f / 255.0f);
            if (!(scale > 0.0f) || !isfinite(scale)) scale = 1.0f;
            zp = __float2int_rn(-mn / scale);
        }
        rec_lo = __float_as_uint(scale), rec_hi = static_cast<uint32_t>(zp);
        const float inv = 1.0f / scale;
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const float r = rintf(x[i] * inv) + static_cast<float>(zp);
            const uint32_t c = i < valid ? static_cast<uint32_t>(fminf(255.0f, fmaxf(0.0f, r))) : 0u;
            w[i >> 2] |= c << (8 * (i & 3));
        }
    } else {
        float amax = 0.0f;
#pragma unroll
        for (int i = 0; i < 16; ++i)
            if (i < valid) amax = fmaxf(amax, fabsf(x[i]));
        amax = group8_max(amax);
        const float scale = amax > 0.0f ? amax / 448.0f : 1.0f;
        rec_lo = __float_as_uint(scale), rec_hi = 0;
        const float inv = 1.0f / scale;
#pragma unroll
        for (int i = 0; i < 16; ++i) w[i >> 2] |= (i < valid ? fp8_encode(x[i] * inv) : 0u) << (8 * (i & 3));
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
}

template <int ALGO> __device__ __forceinline__ void dequantize16(uint4 packed, uint32_t rec_lo, uint32_t rec_hi, float (&out)[16]) {
    const uint32_t w[4] = {packed.x, packed.y, packed.z, packed.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        float y[4];
        dequantize4<ALGO>(w[j], rec_lo, rec_hi, y);
#pragma unroll
        for (int i = 0; i < 4; ++i) out[4 * j + i] = y[i];
    }
}

} // namespace pccl::cuda
