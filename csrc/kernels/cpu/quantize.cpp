#include "quantize.hpp"

#include "../../util/env.hpp"
#include "quantize_fast.hpp"

#include <algorithm>
#include <atomic>
#include <bit>
#include <cmath>
#include <cstring>
#include <limits>

namespace pccl::cpu {

using wire::DType;
using wire::QuantAlgo;
using wire::RedOp;

float fp8e4m3_to_f32(uint8_t v) {
    const uint32_t sign = static_cast<uint32_t>(v & 0x80u) << 24;
    const uint32_t exp = (v >> 3) & 0xfu, man = v & 0x7u;
    if (exp == 0xf && man == 0x7) return std::bit_cast<float>(sign | 0x7fc00000u); // NaN (no inf in e4m3fn)
    if (exp == 0) {
        const float sub = static_cast<float>(man) * (1.0f / 512.0f); // man * 2^-3 * 2^-6
        return (v & 0x80u) ? -sub : sub;
    }
    return std::bit_cast<float>(sign | ((exp + 120u) << 23) | (man << 20));
}

uint8_t f32_to_fp8e4m3(float f) {
    const uint32_t u = std::bit_cast<uint32_t>(f);
    const uint8_t sign = static_cast<uint8_t>((u >> 24) & 0x80u);
    const float a = std::fabs(f);
    if (std::isnan(f)) return static_cast<uint8_t>(sign | 0x7f);
    if (a >= 448.0f) return static_cast<uint8_t>(sign | 0x7e); // saturate (satfinite)
    if (a < 0.0009765625f) return sign;                        // < 2^-10 rounds to zero (half of min subnormal 2^-9)
    int e;
    (void) std::frexp(a, &e); // a = m * 2^e, m in [0.5,1)
    int exp = e - 1;          // a = 1.x * 2^exp
    if (exp < -6) exp = -6;   // subnormal range uses the fixed exponent -6
    const float quantum = std::ldexp(1.0f, exp - 3);
    float steps = a / quantum; // exact: division by a power of two
    float r = std::nearbyint(steps); // default rounding mode is nearest-even
    float val = r * quantum;
    if (val >= 448.0f) return static_cast<uint8_t>(sign | 0x7e);
    if (val == 0.0f) return sign;
    (void) std::frexp(val, &e);
    exp = e - 1;
    if (exp < -6) { // subnormal
        const int man = static_cast<int>(val * 512.0f);
        return static_cast<uint8_t>(sign | man);
    }
    const int man = static_cast<int>((val / std::ldexp(1.0f, exp) - 1.0f) * 8.0f);
    return static_cast<uint8_t>(sign | ((exp + 7) << 3) | man);
}

namespace {

template <typename Q> constexpr double qmax_d() { return static_cast<double>(std::numeric_limits<Q>::max()); }

template <typename Q> Q clamp_to(double v) {
    if (!(v > 0)) {
        if constexpr (std::is_signed_v<Q>) {
            const double lo = static_cast<double>(std::numeric_limits<Q>::lowest());
            if (v <= lo) return std::numeric_limits<Q>::lowest();
            return static_cast<Q>(v);
        }
        return 0;
    }
    if (v >= qmax_d<Q>()) return std::numeric_limits<Q>::max();
    return static_cast<Q>(v);
}

template <typename T> void minmax_scan(const T *x, size_t n, double &mn, double &mx) {
    T lo = std::numeric_limits<T>::max(), hi = std::numeric_limits<T>::lowest();
    for (size_t i = 0; i < n; ++i) {
        lo = x[i] < lo ? x[i] : lo;
        hi = x[i] > hi ? x[i] : hi;
    }
    if (n == 0) lo = hi = 0;
    mn = lo;
    mx = hi;
}

template <typename T, typename Q> void minmax_quant(const T *x, Q *q, size_t n, T mn, T mx) {
    if (mx == mn) {
        if (n) std::memset(q, 0, n * sizeof(Q)); // empty ring chunks come with a null pointer
        return;
    }
    const double inv = 1.0 / static_cast<double>(mx - mn), top = qmax_d<Q>();
    for (size_t i = 0; i < n; ++i) {
        const double v = (static_cast<double>(x[i]) - static_cast<double>(mn)) * inv * top;
        q[i] = clamp_to<Q>(v); // truncating
    }
}

template <typename T> T combine(T a, T b, RedOp op) {
    switch (op) {
    case RedOp::Set: return b;
    case RedOp::Sum:
    case RedOp::Avg: return a + b;
    case RedOp::Prod: return a * b;
    case RedOp::Max: return a > b ? a : b;
    case RedOp::Min: return a < b ? a : b;
    }
    return b;
}

template <typename T, typename Q> void minmax_dequant(T *dst, const Q *q, size_t n, T mn, T mx, RedOp op) {
    const double inv_top = 1.0 / qmax_d<Q>();
    const T range = mx - mn;
    for (size_t i = 0; i < n; ++i) {
        const T v = mn + static_cast<T>(static_cast<double>(q[i]) * inv_top) * range;
        dst[i] = combine(dst[i], v, op);
    }
}

template <typename T, typename Q> void zps_quant(const T *x, Q *q, size_t n, double scale, int64_t zp) {
    const double inv = 1.0 / scale;
    for (size_t i = 0; i < n; ++i) {
        const double v = std::nearbyint(static_cast<double>(x[i]) * inv) + static_cast<double>(zp);
        q[i] = clamp_to<Q>(v);
    }
}

template <typename T, typename Q> void zps_dequant(T *dst, const Q *q, size_t n, double scale, int64_t zp, RedOp op) {
    for (size_t i = 0; i < n; ++i) {
        const T v = static_cast<T>((static_cast<double>(q[i]) - static_cast<double>(zp)) * scale);
        dst[i] = combine(dst[i], v, op);
    }
}

template <typename T, typename Fn> void with_code_type(DType code, Fn fn) {
    switch (code) {
    case DType::U8: fn(uint8_t{}); break;
    case DType::I8: fn(int8_t{}); break;
    case DType::U16: fn(uint16_t{}); break;
    case DType::I16: fn(int16_t{}); break;
    case DType::U32: fn(uint32_t{}); break;
    case DType::I32: fn(int32_t{}); break;
    case DType::U64: fn(uint64_t{}); break;
    case DType::I64: fn(int64_t{}); break;
    default: break;
    }
}

template <typename T> void block_minmax_quant(const T *x, uint8_t *out, size_t n) {
    const size_t nb = (n + kQuantBlock - 1) / kQuantBlock;
    auto *meta = reinterpret_cast<float *>(out);
    uint8_t *codes = out + nb * 2 * sizeof(float);
    for (size_t b = 0; b < nb; ++b) {
        const size_t lo = b * kQuantBlock, hi = std::min(n, lo + kQuantBlock);
        float mn = std::numeric_limits<float>::max(), mx = std::numeric_limits<float>::lowest();
        for (size_t i = lo; i < hi; ++i) {
            const float v = static_cast<float>(x[i]);
            mn = std::min(mn, v);
            mx = std::max(mx, v);
        }
        std::memcpy(&meta[2 * b], &mn, 4);
        std::memcpy(&meta[2 * b + 1], &mx, 4);
        const float inv = mx > mn ? 255.0f / (mx - mn) : 0.0f;
        for (size_t i = lo; i < hi; ++i) {
            const float r = std::nearbyint((static_cast<float>(x[i]) - mn) * inv);
            codes[i] = static_cast<uint8_t>(std::min(255.0f, std::max(0.0f, r)));
        }
    }
}

template <typename T> void block_minmax_dequant(T *dst, const uint8_t *in, size_t n, RedOp op) {
    const size_t nb = (n + kQuantBlock - 1) / kQuantBlock;
    const uint8_t *codes = in + nb * 2 * sizeof(float);
    for (size_t b = 0; b < nb; ++b) {
        float mn, mx;
        std::memcpy(&mn, in + (2 * b) * 4, 4);
        std::memcpy(&mx, in + (2 * b + 1) * 4, 4);
        const float step = (mx - mn) * (1.0f / 255.0f);
        const size_t lo = b * kQuantBlock, hi = std::min(n, lo + kQuantBlock);
        for (size_t i = lo; i < hi; ++i) dst[i] = combine(dst[i], static_cast<T>(std::fmaf(codes[i], step, mn)), op);
    }
}

template <typename T> void block_fp8_quant(const T *x, uint8_t *out, size_t n) {
    const size_t nb = (n + kQuantBlock - 1) / kQuantBlock;
    uint8_t *codes = out + nb * sizeof(float);
    for (size_t b = 0; b < nb; ++b) {
        const size_t lo = b * kQuantBlock, hi = std::min(n, lo + kQuantBlock);
        float amax = 0;
        for (size_t i = lo; i < hi; ++i) amax = std::max(amax, std::fabs(static_cast<float>(x[i])));
        const float scale = amax > 0 ? amax / 448.0f : 1.0f; // x = code * scale
        std::memcpy(out + b * 4, &scale, 4);
        const float inv = 1.0f / scale;
        for (size_t i = lo; i < hi; ++i) codes[i] = f32_to_fp8e4m3(static_cast<float>(x[i]) * inv);
    }
}

template <typename T> void block_fp8_dequant(T *dst, const uint8_t *in, size_t n, RedOp op) {
    const size_t nb = (n + kQuantBlock - 1) / kQuantBlock;
    const uint8_t *codes = in + nb * sizeof(float);
    for (size_t b = 0; b < nb; ++b) {
        float scale;
        std::memcpy(&scale, in + b * 4, 4);
        const size_t lo = b * kQuantBlock, hi = std::min(n, lo + kQuantBlock);
        for (size_t i = lo; i < hi; ++i) dst[i] = combine(dst[i], static_cast<T>(fp8e4m3_to_f32(codes[i]) * scale), op);
    }
}

bool is_int_code(DType t) { return static_cast<uint8_t>(t) <= static_cast<uint8_t>(DType::I64); }
bool is_unsigned_code(DType t) { return t == DType::U8 || t == DType::U16 || t == DType::U32 || t == DType::U64; }

} // namespace

namespace {
std::atomic<int> g_generic{-1};
bool generic_forced() {
    int g = g_generic.load(std::memory_order_relaxed);
    if (g < 0) {
        g = env_flag("PCCL_QUANT_GENERIC", false) ? 1 : 0;
        g_generic.store(g, std::memory_order_relaxed);
    }
    return g == 1;
}
bool use_fast(const QuantParams &p) {
    return !generic_forced() && p.value_dtype == DType::F32 && (p.code_dtype == DType::U8 || p.code_dtype == DType::F8E4M3);
}
} // namespace

void quant_force_generic(bool on) { g_generic.store(on ? 1 : 0, std::memory_order_relaxed); }

bool quant_supported(QuantAlgo algo, DType value_dtype, DType code_dtype) {
    if (value_dtype != DType::F32 && value_dtype != DType::F64) return false;
    switch (algo) {
    case QuantAlgo::None: return true;
    case QuantAlgo::MinMax: return is_unsigned_code(code_dtype);
    case QuantAlgo::ZeroPointScale: return is_int_code(code_dtype);
    case QuantAlgo::BlockMinMax: return code_dtype == DType::U8;
    case QuantAlgo::BlockFp8: return code_dtype == DType::F8E4M3;
    }
    return false;
}

size_t quantized_size(const QuantParams &p, size_t count) {
    const size_t nb = (count + kQuantBlock - 1) / kQuantBlock;
    switch (p.algo) {
    case QuantAlgo::BlockMinMax: return nb * 8 + count;
    case QuantAlgo::BlockFp8: return nb * 4 + count;
    default: return count * wire::dtype_size(p.code_dtype);
    }
}

void quant_analyze(const void *src, size_t count, QuantParams &p) {
    if (p.algo != QuantAlgo::MinMax && p.algo != QuantAlgo::ZeroPointScale) return;
    double mn = 0, mx = 0;
    if (p.value_dtype == DType::F32 && !generic_forced()) {
        float lo, hi;
        fast::extrema_f32(static_cast<const float *>(src), count, lo, hi);
        mn = lo, mx = hi;
    } else if (p.value_dtype == DType::F32)
        minmax_scan(static_cast<const float *>(src), count, mn, mx);
    else
        minmax_scan(static_cast<const double *>(src), count, mn, mx);
    p.min = mn;
    p.max = mx;
    if (p.algo == QuantAlgo::ZeroPointScale) {
        if (mx == mn) {
            // constant chunk: q = 1, scale = the value itself (q = 0 for an all-zero chunk)
            p.scale = mn != 0 ? mn : 1.0;
            p.zero_point = 0;
            return;
        }
        double qlo = 0, qhi = 255;
        with_code_type<float>(p.code_dtype, [&](auto tag) {
            using Q = decltype(tag);
            qlo = static_cast<double>(std::numeric_limits<Q>::lowest());
            qhi = static_cast<double>(std::numeric_limits<Q>::max());
        });
        p.scale = (mx - mn) / (qhi - qlo);
        if (p.value_dtype == DType::F32) p.scale = static_cast<double>(static_cast<float>(p.scale));
        if (!(p.scale > 0) || !std::isfinite(p.scale)) p.scale = 1.0;
        p.zero_point = static_cast<int64_t>(std::nearbyint(qlo - mn / p.scale));
    }
}

void quantize(const void *src, void *dst, size_t count, const QuantParams &p) {
    const bool f32 = p.value_dtype == DType::F32;
    if (use_fast(p)) {
        const auto *x = static_cast<const float *>(src);
        auto *q = static_cast<uint8_t *>(dst);
        switch (p.algo) {
        case QuantAlgo::MinMax: fast::minmax_quant_u8(x, q, count, static_cast<float>(p.min), static_cast<float>(p.max)); return;
        case QuantAlgo::ZeroPointScale: fast::zps_quant_u8(x, q, count, p.scale, p.zero_point); return;
        case QuantAlgo::BlockMinMax: fast::block_minmax_quant(x, q, count); return;
        case QuantAlgo::BlockFp8: fast::block_fp8_quant(x, q, count); return;
        default: break;
        }
    }
    switch (p.algo) {
    case QuantAlgo::None: std::memcpy(dst, src, count * wire::dtype_size(p.value_dtype)); break;
    case QuantAlgo::MinMax:
        with_code_type<float>(p.code_dtype, [&](auto tag) {
            using Q = decltype(tag);
            if (f32)
                minmax_quant(static_cast<const float *>(src), static_cast<Q *>(dst), count, static_cast<float>(p.min),
                             static_cast<float>(p.max));
            else
                minmax_quant(static_cast<const double *>(src), static_cast<Q *>(dst), count, p.min, p.max);
        });
        break;
    case QuantAlgo::ZeroPointScale:
        with_code_type<float>(p.code_dtype, [&](auto tag) {
            using Q = decltype(tag);
            if (f32)
                zps_quant(static_cast<const float *>(src), static_cast<Q *>(dst), count, p.scale, p.zero_point);
            else
                zps_quant(static_cast<const double *>(src), static_cast<Q *>(dst), count, p.scale, p.zero_point);
        });
        break;
    case QuantAlgo::BlockMinMax:
        f32 ? block_minmax_quant(static_cast<const float *>(src), static_cast<uint8_t *>(dst), count)
            : block_minmax_quant(static_cast<const double *>(src), static_cast<uint8_t *>(dst), count);
        break;
    case QuantAlgo::BlockFp8:
        f32 ? block_fp8_quant(static_cast<const float *>(src), static_cast<uint8_t *>(dst), count)
            : block_fp8_quant(static_cast<const double *>(src), static_cast<uint8_t *>(dst), count);
        break;
    }
}

void dequantize_reduce(void *dst, const void *q, size_t count, const QuantParams &p, RedOp op) {
    const bool f32 = p.value_dtype == DType::F32;
    const bool add = op == RedOp::Sum || op == RedOp::Avg, set = op == RedOp::Set;
    if (use_fast(p) && (add || set)) { // the other operators are rare with quantisation and keep the generic loops
        auto *d = static_cast<float *>(dst);
        const auto *c = static_cast<const uint8_t *>(q);
        switch (p.algo) {
        case QuantAlgo::MinMax:
            (set ? fast::minmax_dequant_u8_set : fast::minmax_dequant_u8_add)(d, c, count, static_cast<float>(p.min), static_cast<float>(p.max));
            return;
        case QuantAlgo::ZeroPointScale: (set ? fast::zps_dequant_u8_set : fast::zps_dequant_u8_add)(d, c, count, p.scale, p.zero_point); return;
        case QuantAlgo::BlockMinMax:
            if (fast::cpu_has_fma()) {
                (set ? fast::block_minmax_dequant_set_fma : fast::block_minmax_dequant_add_fma)(d, c, count);
                return;
            }
            break;
        case QuantAlgo::BlockFp8: (set ? fast::block_fp8_dequant_set : fast::block_fp8_dequant_add)(d, c, count); return;
        default: break;
        }
    }
    switch (p.algo) {
    case QuantAlgo::None: break;
    case QuantAlgo::MinMax:
        with_code_type<float>(p.code_dtype, [&](auto tag) {
            using Q = decltype(tag);
            if (f32)
                minmax_dequant(static_cast<float *>(dst), static_cast<const Q *>(q), count, static_cast<float>(p.min),
                               static_cast<float>(p.max), op);
            else
                minmax_dequant(static_cast<double *>(dst), static_cast<const Q *>(q), count, p.min, p.max, op);
        });
        break;
    case QuantAlgo::ZeroPointScale:
        with_code_type<float>(p.code_dtype, [&](auto tag) {
            using Q = decltype(tag);
            if (f32)
                zps_dequant(static_cast<float *>(dst), static_cast<const Q *>(q), count, p.scale, p.zero_point, op);
            else
                zps_dequant(static_cast<double *>(dst), static_cast<const Q *>(q), count, p.scale, p.zero_point, op);
        });
        break;
    case QuantAlgo::BlockMinMax:
        f32 ? block_minmax_dequant(static_cast<float *>(dst), static_cast<const uint8_t *>(q), count, op)
            : block_minmax_dequant(static_cast<double *>(dst), static_cast<const uint8_t *>(q), count, op);
        break;
    case QuantAlgo::BlockFp8:
        f32 ? block_fp8_dequant(static_cast<float *>(dst), static_cast<const uint8_t *>(q), count, op)
            : block_fp8_dequant(static_cast<double *>(dst), static_cast<const uint8_t *>(q), count, op);
        break;
    }
}

// Floating point metadata travels in host byte order, integers big-endian (reference:
// ccoip/internal/quantize.hpp:57-69).
wire::P2PDequantMeta to_wire(const QuantParams &p, uint64_t tag) {
    wire::P2PDequantMeta m;
    m.tag = tag;
    auto put_float = [&](std::vector<uint8_t> &blob, double v) {
        if (p.value_dtype == DType::F32) {
            const float f = static_cast<float>(v);
            blob.resize(4);
            std::memcpy(blob.data(), &f, 4);
        } else {
            blob.resize(8);
            std::memcpy(blob.data(), &v, 8);
        }
    };
    if (p.algo == QuantAlgo::ZeroPointScale) {
        m.meta_type = wire::DequantMetaType::ZeroPointScale;
        m.dtype_a = DType::I64;
        m.blob_a.resize(8);
        uint64_t z = static_cast<uint64_t>(p.zero_point);
        for (int i = 7; i >= 0; --i) {
            m.blob_a[i] = static_cast<uint8_t>(z);
            z >>= 8;
        }
        m.dtype_b = p.value_dtype;
        put_float(m.blob_b, p.scale);
    } else {
        m.meta_type = wire::DequantMetaType::MinMax;
        m.dtype_a = m.dtype_b = p.value_dtype;
        put_float(m.blob_a, p.min);
        put_float(m.blob_b, p.max);
    }
    return m;
}

bool from_wire(const wire::P2PDequantMeta &m, QuantParams &p) {
    auto get_float = [&](const std::vector<uint8_t> &blob, DType t, double &out) {
        if (t == DType::F32 && blob.size() == 4) {
            float f;
            std::memcpy(&f, blob.data(), 4);
            out = f;
            return true;
        }
        if (t == DType::F64 && blob.size() == 8) {
            std::memcpy(&out, blob.data(), 8);
            return true;
        }
        return false;
    };
    if (m.meta_type == wire::DequantMetaType::ZeroPointScale) {
        if (m.dtype_a != DType::I64 || m.blob_a.size() != 8) return false;
        uint64_t z = 0;
        for (int i = 0; i < 8; ++i) z = (z << 8) | m.blob_a[i];
        p.zero_point = static_cast<int64_t>(z);
        return get_float(m.blob_b, m.dtype_b, p.scale);
    }
    return get_float(m.blob_a, m.dtype_a, p.min) && get_float(m.blob_b, m.dtype_b, p.max);
}

} // namespace pccl::cpu
