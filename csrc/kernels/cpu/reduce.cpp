#include "reduce.hpp"

#include <algorithm>
#include <bit>
#include <cstring>
#include <thread>
#include <vector>

#include "half.hpp"

namespace pccl::cpu {

using wire::DType;
using wire::RedOp;

namespace {

template <typename T, typename F> void apply(T *__restrict dst, const T *__restrict src, size_t n, F f) {
#pragma GCC ivdep
    for (size_t i = 0; i < n; ++i) dst[i] = f(dst[i], src[i]);
}

template <typename T> void reduce_typed(void *dstv, const void *srcv, size_t n, RedOp op) {
    T *dst = static_cast<T *>(dstv);
    const T *src = static_cast<const T *>(srcv);
    switch (op) {
    case RedOp::Set: std::memcpy(dst, src, n * sizeof(T)); break;
    case RedOp::Sum:
    case RedOp::Avg: apply(dst, src, n, [](T a, T b) { return static_cast<T>(a + b); }); break;
    case RedOp::Prod: apply(dst, src, n, [](T a, T b) { return static_cast<T>(a * b); }); break;
    case RedOp::Max: apply(dst, src, n, [](T a, T b) { return a > b ? a : b; }); break;
    case RedOp::Min: apply(dst, src, n, [](T a, T b) { return a < b ? a : b; }); break;
    }
}

// 16-bit floats: operate in fp32, round once per operation.  half.hpp defines the conversions; the loops below produce the
// same bits faster -- bfloat16 with a branch-free rounding the compiler vectorises (one clone per instruction set), float16
// with the F16C conversion instructions where the CPU has them (their NaNs are made canonical like the software path's).
#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__) && !defined(__SANITIZE_ADDRESS__) && !defined(__SANITIZE_THREAD__)
#define PCCL_CLONES __attribute__((target_clones("arch=x86-64-v4", "arch=x86-64-v3", "default")))
#define PCCL_HAVE_F16C_PATH 1
#else
#define PCCL_CLONES
#define PCCL_HAVE_F16C_PATH 0
#endif

enum class HalfOp { Add, Mul, Max, Min, Div };

template <HalfOp OP> __attribute__((always_inline)) inline float half_op(float a, float b) {
    if constexpr (OP == HalfOp::Add) return a + b;
    if constexpr (OP == HalfOp::Mul) return a * b;
    if constexpr (OP == HalfOp::Max) return a > b ? a : b;
    if constexpr (OP == HalfOp::Min) return a < b ? a : b;
    return a / b;
}

__attribute__((always_inline)) inline uint16_t bf16_round(float f) { // == f32_to_bf16, without the branch
    const uint32_t u = std::bit_cast<uint32_t>(f);
    const uint32_t rounded = (u + 0x7fffu + ((u >> 16) & 1u)) >> 16, nan = (u >> 16) | 0x40u;
    return static_cast<uint16_t>((u & 0x7fffffffu) > 0x7f800000u ? nan : rounded);
}

template <HalfOp OP> __attribute__((always_inline)) inline void bf16_loop(uint16_t *__restrict dst, const uint16_t *__restrict src, size_t n, float scalar) {
    for (size_t i = 0; i < n; ++i) {
        const float a = std::bit_cast<float>(static_cast<uint32_t>(dst[i]) << 16);
        const float b = OP == HalfOp::Div ? scalar : std::bit_cast<float>(static_cast<uint32_t>(src[i]) << 16);
        dst[i] = bf16_round(half_op<OP>(a, b));
    }
}
PCCL_CLONES void bf16_add(uint16_t *dst, const uint16_t *src, size_t n) { bf16_loop<HalfOp::Add>(dst, src, n, 0); }
PCCL_CLONES void bf16_mul(uint16_t *dst, const uint16_t *src, size_t n) { bf16_loop<HalfOp::Mul>(dst, src, n, 0); }
PCCL_CLONES void bf16_max(uint16_t *dst, const uint16_t *src, size_t n) { bf16_loop<HalfOp::Max>(dst, src, n, 0); }
PCCL_CLONES void bf16_min(uint16_t *dst, const uint16_t *src, size_t n) { bf16_loop<HalfOp::Min>(dst, src, n, 0); }
PCCL_CLONES void bf16_div(uint16_t *dst, size_t n, float by) {
    for (size_t i = 0; i < n; ++i) dst[i] = bf16_round(std::bit_cast<float>(static_cast<uint32_t>(dst[i]) << 16) / by);
}

template <HalfOp OP> void f16_software(uint16_t *dst, const uint16_t *src, size_t n, float scalar) {
    for (size_t i = 0; i < n; ++i) dst[i] = f32_to_f16(half_op<OP>(f16_to_f32(dst[i]), OP == HalfOp::Div ? scalar : f16_to_f32(src[i])));
}

#if PCCL_HAVE_F16C_PATH
} // namespace
} // namespace pccl::cpu
#include <immintrin.h>
namespace pccl::cpu {
namespace {
template <HalfOp OP> __attribute__((target("avx2,f16c"))) void f16_hardware(uint16_t *dst, const uint16_t *src, size_t n, float scalar) {
    const __m256 s = _mm256_set1_ps(scalar);
    const __m128i mag_mask = _mm_set1_epi16(0x7fff), inf = _mm_set1_epi16(0x7c00), sign_mask = _mm_set1_epi16(static_cast<short>(0x8000)),
                  quiet = _mm_set1_epi16(0x7e00);
    size_t i = 0;
    for (; i + 8 <= n; i += 8) {
        const __m256 a = _mm256_cvtph_ps(_mm_loadu_si128(reinterpret_cast<const __m128i *>(dst + i)));
        const __m256 b = OP == HalfOp::Div ? s : _mm256_cvtph_ps(_mm_loadu_si128(reinterpret_cast<const __m128i *>(src + i)));
        __m256 r;
        if constexpr (OP == HalfOp::Add) r = _mm256_add_ps(a, b);
        else if constexpr (OP == HalfOp::Mul) r = _mm256_mul_ps(a, b);
        else if constexpr (OP == HalfOp::Max) r = _mm256_blendv_ps(b, a, _mm256_cmp_ps(a, b, _CMP_GT_OQ));
        else if constexpr (OP == HalfOp::Min) r = _mm256_blendv_ps(b, a, _mm256_cmp_ps(a, b, _CMP_LT_OQ));
        else r = _mm256_div_ps(a, b);
        __m128i h = _mm256_cvtps_ph(r, _MM_FROUND_TO_NEAREST_INT | _MM_FROUND_NO_EXC);
        const __m128i is_nan = _mm_cmpgt_epi16(_mm_and_si128(h, mag_mask), inf); // magnitudes are non-negative as int16
        h = _mm_blendv_epi8(h, _mm_or_si128(_mm_and_si128(h, sign_mask), quiet), is_nan);
        _mm_storeu_si128(reinterpret_cast<__m128i *>(dst + i), h);
    }
    f16_software<OP>(dst + i, src + i, n - i, scalar);
}
bool have_f16c() {
    static const bool yes = __builtin_cpu_supports("avx2") && __builtin_cpu_supports("f16c");
    return yes;
}
#else
bool have_f16c() { return false; }
template <HalfOp OP> void f16_hardware(uint16_t *dst, const uint16_t *src, size_t n, float scalar) { f16_software<OP>(dst, src, n, scalar); }
#endif

template <HalfOp OP> void f16_apply(uint16_t *dst, const uint16_t *src, size_t n, float scalar = 0) {
    have_f16c() ? f16_hardware<OP>(dst, src, n, scalar) : f16_software<OP>(dst, src, n, scalar);
}

void reduce_f16(void *dstv, const void *srcv, size_t n, RedOp op) {
    auto *dst = static_cast<uint16_t *>(dstv);
    const auto *src = static_cast<const uint16_t *>(srcv);
    switch (op) {
    case RedOp::Set: std::memcpy(dst, src, n * 2); break;
    case RedOp::Sum:
    case RedOp::Avg: f16_apply<HalfOp::Add>(dst, src, n); break;
    case RedOp::Prod: f16_apply<HalfOp::Mul>(dst, src, n); break;
    case RedOp::Max: f16_apply<HalfOp::Max>(dst, src, n); break;
    case RedOp::Min: f16_apply<HalfOp::Min>(dst, src, n); break;
    }
}

void reduce_bf16(void *dstv, const void *srcv, size_t n, RedOp op) {
    auto *dst = static_cast<uint16_t *>(dstv);
    const auto *src = static_cast<const uint16_t *>(srcv);
    switch (op) {
    case RedOp::Set: std::memcpy(dst, src, n * 2); break;
    case RedOp::Sum:
    case RedOp::Avg: bf16_add(dst, src, n); break;
    case RedOp::Prod: bf16_mul(dst, src, n); break;
    case RedOp::Max: bf16_max(dst, src, n); break;
    case RedOp::Min: bf16_min(dst, src, n); break;
    }
}

template <typename T> void avg_typed(void *dstv, size_t n, uint64_t world) {
    T *dst = static_cast<T *>(dstv);
    const T w = static_cast<T>(world);
    for (size_t i = 0; i < n; ++i) dst[i] = static_cast<T>(dst[i] / w);
}

} // namespace

bool dtype_supported(DType t) { return static_cast<uint8_t>(t) <= static_cast<uint8_t>(DType::F64); }

void reduce_inplace(void *dst, const void *src, size_t count, DType dtype, RedOp op) {
    switch (dtype) {
    case DType::U8: reduce_typed<uint8_t>(dst, src, count, op); break;
    case DType::I8: reduce_typed<int8_t>(dst, src, count, op); break;
    case DType::U16: reduce_typed<uint16_t>(dst, src, count, op); break;
    case DType::I16: reduce_typed<int16_t>(dst, src, count, op); break;
    case DType::U32: reduce_typed<uint32_t>(dst, src, count, op); break;
    case DType::I32: reduce_typed<int32_t>(dst, src, count, op); break;
    case DType::U64: reduce_typed<uint64_t>(dst, src, count, op); break;
    case DType::I64: reduce_typed<int64_t>(dst, src, count, op); break;
    case DType::F32: reduce_typed<float>(dst, src, count, op); break;
    case DType::F64: reduce_typed<double>(dst, src, count, op); break;
    case DType::F16: reduce_f16(dst, src, count, op); break;
    case DType::BF16: reduce_bf16(dst, src, count, op); break;
    case DType::F8E4M3: break;
    }
}

namespace {
void finalize_avg_serial(void *dst, size_t count, DType dtype, uint64_t world);
} // namespace

// The division is one more pass over the whole result; large results are cut into a few ranges that are divided at once.
void finalize_avg(void *dst, size_t count, DType dtype, uint64_t world) {
    if (world <= 1) return;
    const size_t esz = wire::dtype_size(dtype);
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    const size_t crew = std::min<size_t>({4, hw / 2, count * esz / (16u << 20)});
    if (crew < 2) return finalize_avg_serial(dst, count, dtype, world);
    const size_t per = ((count + crew - 1) / crew + 63) / 64 * 64;
    std::vector<std::thread> helpers;
    for (size_t t = 1; t < crew; ++t) {
        const size_t lo = std::min(count, t * per), hi = std::min(count, lo + per);
        if (lo < hi) helpers.emplace_back([=] { finalize_avg_serial(static_cast<uint8_t *>(dst) + lo * esz, hi - lo, dtype, world); });
    }
    finalize_avg_serial(dst, std::min(count, per), dtype, world);
    for (auto &h : helpers) h.join();
}

namespace {
void finalize_avg_serial(void *dst, size_t count, DType dtype, uint64_t world) {
    switch (dtype) {
    case DType::U8: avg_typed<uint8_t>(dst, count, world); break;
    case DType::I8: avg_typed<int8_t>(dst, count, world); break;
    case DType::U16: avg_typed<uint16_t>(dst, count, world); break;
    case DType::I16: avg_typed<int16_t>(dst, count, world); break;
    case DType::U32: avg_typed<uint32_t>(dst, count, world); break;
    case DType::I32: avg_typed<int32_t>(dst, count, world); break;
    case DType::U64: avg_typed<uint64_t>(dst, count, world); break;
    case DType::I64: avg_typed<int64_t>(dst, count, world); break;
    case DType::F32: avg_typed<float>(dst, count, world); break;
    case DType::F64: avg_typed<double>(dst, count, world); break;
    case DType::F16: f16_apply<HalfOp::Div>(static_cast<uint16_t *>(dst), static_cast<const uint16_t *>(dst), count, static_cast<float>(world)); break;
    case DType::BF16: bf16_div(static_cast<uint16_t *>(dst), count, static_cast<float>(world)); break;
    case DType::F8E4M3: break;
    }
}
} // namespace

} // namespace pccl::cpu
