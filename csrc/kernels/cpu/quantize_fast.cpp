// Fast paths of the host quantisers for the combination that matters on the wire: float32 values, 8-bit codes.
//
// The generic templates in quantize.cpp define the arithmetic (min-max: the reference's formula, in double, truncating --
// ccoip/src/cpp/quantize_kernels.cpp:24-58; zero-point/scale: round to nearest; the two block formats: this library's
// own).  They run at 0.1-2.4 GB/s of fp32 input because of branches inside the loops (clamping, the operator switch, the
// scalar fp8 conversion), which is less than a 10 Gbit/s link carries once the data is a quarter of its size.  The loops here
// compute EXACTLY the same values -- tests/cpp/test_units.cpp compares them bit for bit -- but are written so that the compiler
// vectorises them (no branches, no calls, operator hoisted out), and are compiled once per instruction set (function
// multi-versioning, resolved when the library is loaded).  No contraction into fused multiply-adds is allowed anywhere
// (-ffp-contract=off): a*b+c must round twice wherever the generic code rounds twice.
#include "quantize_fast.hpp"

#include <bit>
#include <cmath>
#include <cstring>
#include <limits>

#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__) && !defined(__SANITIZE_ADDRESS__) && !defined(__SANITIZE_THREAD__)
#define PCCL_CLONES __attribute__((target_clones("arch=x86-64-v4", "arch=x86-64-v3", "default")))
#else
#define PCCL_CLONES
#endif

namespace pccl::cpu::fast {

using wire::RedOp;

// ---- extrema --------------------------------------------------------------------------------------------------------
// Independent lanes instead of one running pair: the inner loop is element-wise, so it becomes packed min/max.  `v < lo ?
// v : lo` is what minps computes; a NaN never wins a comparison, so NaNs are skipped here exactly as in the sequential scan.
PCCL_CLONES void extrema_f32(const float *x, size_t n, float &lo_out, float &hi_out) {
    constexpr int kLanes = 64;
    float lo[kLanes], hi[kLanes];
    for (int j = 0; j < kLanes; ++j) lo[j] = std::numeric_limits<float>::max(), hi[j] = std::numeric_limits<float>::lowest();
    size_t i = 0;
    for (; i + kLanes <= n; i += kLanes)
        for (int j = 0; j < kLanes; ++j) {
            const float v = x[i + j];
            lo[j] = v < lo[j] ? v : lo[j];
            hi[j] = v > hi[j] ? v : hi[j];
        }
    float l = std::numeric_limits<float>::max(), h = std::numeric_limits<float>::lowest();
    for (; i < n; ++i) {
        l = x[i] < l ? x[i] : l;
        h = x[i] > h ? x[i] : h;
    }
    for (int j = 0; j < kLanes; ++j) {
        l = lo[j] < l ? lo[j] : l;
        h = hi[j] > h ? hi[j] : h;
    }
    if (n == 0) l = h = 0;
    lo_out = l, hi_out = h;
}

// ---- min-max (one range per chunk) -----------------------------------------------------------------------------------
PCCL_CLONES void minmax_quant_u8(const float *x, uint8_t *q, size_t n, float mn, float mx) {
    if (mx == mn) {
        if (n) std::memset(q, 0, n);
        return;
    }
    const double inv = 1.0 / static_cast<double>(mx - mn), lo = static_cast<double>(mn);
    for (size_t i = 0; i < n; ++i) {
        double v = (static_cast<double>(x[i]) - lo) * inv * 255.0;
        v = v > 0.0 ? v : 0.0; // also catches NaN, like the generic clamp
        v = v < 255.0 ? v : 255.0;
        q[i] = static_cast<uint8_t>(static_cast<int32_t>(v)); // truncating
    }
}

template <bool SET> __attribute__((always_inline)) static inline void minmax_dequant_u8_body(float *dst, const uint8_t *q, size_t n, float mn, float mx) {
    const double inv_top = 1.0 / 255.0;
    const float range = mx - mn;
    for (size_t i = 0; i < n; ++i) {
        const float v = mn + static_cast<float>(static_cast<double>(q[i]) * inv_top) * range;
        dst[i] = SET ? v : dst[i] + v;
    }
}
PCCL_CLONES void minmax_dequant_u8_set(float *dst, const uint8_t *q, size_t n, float mn, float mx) { minmax_dequant_u8_body<true>(dst, q, n, mn, mx); }
PCCL_CLONES void minmax_dequant_u8_add(float *dst, const uint8_t *q, size_t n, float mn, float mx) { minmax_dequant_u8_body<false>(dst, q, n, mn, mx); }

// ---- zero point / scale ----------------------------------------------------------------------------------------------
PCCL_CLONES void zps_quant_u8(const float *x, uint8_t *q, size_t n, double scale, int64_t zp) {
    const double inv = 1.0 / scale, z = static_cast<double>(zp);
    for (size_t i = 0; i < n; ++i) {
        double v = std::nearbyint(static_cast<double>(x[i]) * inv) + z;
        v = v > 0.0 ? v : 0.0;
        v = v < 255.0 ? v : 255.0;
        q[i] = static_cast<uint8_t>(static_cast<int32_t>(v));
    }
}

template <bool SET> __attribute__((always_inline)) static inline void zps_dequant_u8_body(float *dst, const uint8_t *q, size_t n, double scale, int64_t zp) {
    const double z = static_cast<double>(zp);
    for (size_t i = 0; i < n; ++i) {
        const float v = static_cast<float>((static_cast<double>(q[i]) - z) * scale);
        dst[i] = SET ? v : dst[i] + v;
    }
}
PCCL_CLONES void zps_dequant_u8_set(float *dst, const uint8_t *q, size_t n, double scale, int64_t zp) { zps_dequant_u8_body<true>(dst, q, n, scale, zp); }
PCCL_CLONES void zps_dequant_u8_add(float *dst, const uint8_t *q, size_t n, double scale, int64_t zp) { zps_dequant_u8_body<false>(dst, q, n, scale, zp); }

// ---- uint8 with a range per 128 elements ------------------------------------------------------------------------------
PCCL_CLONES void block_minmax_quant(const float *x, uint8_t *out, size_t n) {
    const size_t nb = (n + kQuantBlock - 1) / kQuantBlock;
    uint8_t *codes = out + nb * 2 * sizeof(float);
    for (size_t b = 0; b < nb; ++b) {
        const size_t lo = b * kQuantBlock, len = (n - lo < kQuantBlock) ? n - lo : kQuantBlock;
        const float *xb = x + lo;
        float mn = std::numeric_limits<float>::max(), mx = std::numeric_limits<float>::lowest();
        if (len == kQuantBlock) {
            float l[16], h[16];
            for (int j = 0; j < 16; ++j) l[j] = mn, h[j] = mx;
            for (size_t i = 0; i < kQuantBlock; i += 16)
                for (int j = 0; j < 16; ++j) {
                    const float v = xb[i + j];
                    l[j] = v < l[j] ? v : l[j]; // std::min(mn, v) of the generic loop keeps mn on NaN; so does this
                    h[j] = v > h[j] ? v : h[j];
                }
            for (int j = 0; j < 16; ++j) {
                mn = l[j] < mn ? l[j] : mn;
                mx = h[j] > mx ? h[j] : mx;
            }
        } else {
            for (size_t i = 0; i < len; ++i) {
                mn = xb[i] < mn ? xb[i] : mn;
                mx = xb[i] > mx ? xb[i] : mx;
            }
        }
        std::memcpy(out + (2 * b) * 4, &mn, 4);
        std::memcpy(out + (2 * b + 1) * 4, &mx, 4);
        const float inv = mx > mn ? 255.0f / (mx - mn) : 0.0f;
        uint8_t *cb = codes + lo;
        for (size_t i = 0; i < len; ++i) {
            float r = std::nearbyint((xb[i] - mn) * inv);
            r = r > 0.0f ? r : 0.0f; // generic: min(255, max(0, r)) with std::max(0, NaN) = 0
            r = r < 255.0f ? r : 255.0f;
            cb[i] = static_cast<uint8_t>(static_cast<int32_t>(r));
        }
    }
}

template <bool SET> __attribute__((always_inline)) static inline void block_minmax_dequant_body(float *dst, const uint8_t *in, size_t n) {
    const size_t nb = (n + kQuantBlock - 1) / kQuantBlock;
    const uint8_t *codes = in + nb * 2 * sizeof(float);
    for (size_t b = 0; b < nb; ++b) {
        float mn, mx;
        std::memcpy(&mn, in + (2 * b) * 4, 4);
        std::memcpy(&mx, in + (2 * b + 1) * 4, 4);
        const float step = (mx - mn) * (1.0f / 255.0f);
        const size_t lo = b * kQuantBlock, len = (n - lo < kQuantBlock) ? n - lo : kQuantBlock;
        for (size_t i = 0; i < len; ++i) {
            const float v = __builtin_fmaf(static_cast<float>(codes[lo + i]), step, mn); // one rounding, as the generic std::fmaf
            dst[lo + i] = SET ? v : dst[lo + i] + v;
        }
    }
}
// without FMA hardware fmaf is a library call per element: these two have no "default" vector clone worth having, the
// dispatcher only uses them when the CPU has FMA
__attribute__((target("avx2,fma"))) void block_minmax_dequant_set_fma(float *dst, const uint8_t *in, size_t n) { block_minmax_dequant_body<true>(dst, in, n); }
__attribute__((target("avx2,fma"))) void block_minmax_dequant_add_fma(float *dst, const uint8_t *in, size_t n) { block_minmax_dequant_body<false>(dst, in, n); }

// ---- fp8 e4m3 with a scale per 128 elements ----------------------------------------------------------------------------
// Round to nearest even at the granularity of the target exponent by adding and subtracting 1.5 * 2^(23 + exponent of the
// quantum): the sum has no room for bits below the quantum, the hardware rounds them away (nearest even), the subtraction is
// exact.  Values below 2^-6 share the subnormal quantum 2^-9.  Bit-identical to f32_to_fp8e4m3() in quantize.cpp.
__attribute__((always_inline)) static inline uint8_t encode_e4m3(float f) {
    const uint32_t u = std::bit_cast<uint32_t>(f);
    const uint32_t sign = (u >> 24) & 0x80u;
    float a = std::bit_cast<float>(u & 0x7fffffffu);
    const bool is_nan = a != a;
    a = a < 448.0f ? a : 448.0f; // saturate; a NaN becomes 448 here and is replaced below
    uint32_t ebits = std::bit_cast<uint32_t>(a) >> 23;
    ebits = ebits > 121u ? ebits : 121u; // exponent -6 is the smallest: subnormals
    const float magic = std::bit_cast<float>(((ebits + 20u) << 23) | 0x400000u); // 1.5 * 2^(e - 3 + 23)
    const float val = (a + magic) - magic;
    const uint32_t vb = std::bit_cast<uint32_t>(val);
    const uint32_t normal = (vb >> 20) - 960u;               // (exponent field - 120) << 3 | top three mantissa bits
    const uint32_t subnormal = static_cast<uint32_t>(static_cast<int32_t>(val * 512.0f));
    uint32_t code = val < 0.015625f ? subnormal : normal;
    code = is_nan ? 0x7fu : code;
    return static_cast<uint8_t>(sign | code);
}

__attribute__((always_inline)) static inline float decode_e4m3(uint8_t c) {
    const uint32_t mag = c & 0x7fu;
    // exponent and mantissa moved into float position give 2^-120 times the value (subnormals included: the float is then
    // subnormal too and the product is exact)
    float v = std::bit_cast<float>(mag << 20) * 0x1p120f;
    v = mag == 0x7fu ? std::numeric_limits<float>::quiet_NaN() : v;
    return std::bit_cast<float>(std::bit_cast<uint32_t>(v) | (static_cast<uint32_t>(c & 0x80u) << 24));
}

uint8_t f32_to_fp8e4m3_branchless(float f) { return encode_e4m3(f); }
float fp8e4m3_to_f32_branchless(uint8_t c) { return decode_e4m3(c); }

PCCL_CLONES void block_fp8_quant(const float *x, uint8_t *out, size_t n) {
    const size_t nb = (n + kQuantBlock - 1) / kQuantBlock;
    uint8_t *codes = out + nb * sizeof(float);
    for (size_t b = 0; b < nb; ++b) {
        const size_t lo = b * kQuantBlock, len = (n - lo < kQuantBlock) ? n - lo : kQuantBlock;
        const float *xb = x + lo;
        float amax = 0;
        if (len == kQuantBlock) {
            float m[16];
            for (int j = 0; j < 16; ++j) m[j] = 0;
            for (size_t i = 0; i < kQuantBlock; i += 16)
                for (int j = 0; j < 16; ++j) {
                    const float a = std::bit_cast<float>(std::bit_cast<uint32_t>(xb[i + j]) & 0x7fffffffu);
                    m[j] = m[j] < a ? a : m[j]; // std::max(amax, a)
                }
            for (int j = 0; j < 16; ++j) amax = amax < m[j] ? m[j] : amax;
        } else {
            for (size_t i = 0; i < len; ++i) {
                const float a = std::fabs(xb[i]);
                amax = amax < a ? a : amax;
            }
        }
        const float scale = amax > 0 ? amax / 448.0f : 1.0f;
        std::memcpy(out + b * 4, &scale, 4);
        const float inv = 1.0f / scale;
        uint8_t *cb = codes + lo;
        for (size_t i = 0; i < len; ++i) cb[i] = encode_e4m3(xb[i] * inv);
    }
}

template <bool SET> __attribute__((always_inline)) static inline void block_fp8_dequant_body(float *dst, const uint8_t *in, size_t n) {
    const size_t nb = (n + kQuantBlock - 1) / kQuantBlock;
    const uint8_t *codes = in + nb * sizeof(float);
    for (size_t b = 0; b < nb; ++b) {
        float scale;
        std::memcpy(&scale, in + b * 4, 4);
        const size_t lo = b * kQuantBlock, len = (n - lo < kQuantBlock) ? n - lo : kQuantBlock;
        for (size_t i = 0; i < len; ++i) {
            const float v = decode_e4m3(codes[lo + i]) * scale;
            dst[lo + i] = SET ? v : dst[lo + i] + v;
        }
    }
}
PCCL_CLONES void block_fp8_dequant_set(float *dst, const uint8_t *in, size_t n) { block_fp8_dequant_body<true>(dst, in, n); }
PCCL_CLONES void block_fp8_dequant_add(float *dst, const uint8_t *in, size_t n) { block_fp8_dequant_body<false>(dst, in, n); }

bool cpu_has_fma() {
#if defined(__x86_64__) && defined(__GNUC__)
    static const bool yes = __builtin_cpu_supports("avx2") && __builtin_cpu_supports("fma");
    return yes;
#else
    return false;
#endif
}

} // namespace pccl::cpu::fast
