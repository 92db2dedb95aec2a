#include "reduce.hpp"

#include <algorithm>
#include <cstring>

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

// 16-bit floats: operate in fp32, round once per operation.
template <float (*Load)(uint16_t), uint16_t (*Store)(float)>
void reduce_half(void *dstv, const void *srcv, size_t n, RedOp op) {
    auto *dst = static_cast<uint16_t *>(dstv);
    const auto *src = static_cast<const uint16_t *>(srcv);
    if (op == RedOp::Set) {
        std::memcpy(dst, src, n * 2);
        return;
    }
    for (size_t i = 0; i < n; ++i) {
        const float a = Load(dst[i]), b = Load(src[i]);
        float r;
        switch (op) {
        case RedOp::Sum:
        case RedOp::Avg: r = a + b; break;
        case RedOp::Prod: r = a * b; break;
        case RedOp::Max: r = a > b ? a : b; break;
        default: r = a < b ? a : b; break;
        }
        dst[i] = Store(r);
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
    case DType::F16: reduce_half<f16_to_f32, f32_to_f16>(dst, src, count, op); break;
    case DType::BF16: reduce_half<bf16_to_f32, f32_to_bf16>(dst, src, count, op); break;
    case DType::F8E4M3: break;
    }
}

void finalize_avg(void *dst, size_t count, DType dtype, uint64_t world) {
    if (world <= 1) return;
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
    case DType::F16: {
        auto *p = static_cast<uint16_t *>(dst);
        for (size_t i = 0; i < count; ++i) p[i] = f32_to_f16(f16_to_f32(p[i]) / static_cast<float>(world));
        break;
    }
    case DType::BF16: {
        auto *p = static_cast<uint16_t *>(dst);
        for (size_t i = 0; i < count; ++i) p[i] = f32_to_bf16(bf16_to_f32(p[i]) / static_cast<float>(world));
        break;
    }
    case DType::F8E4M3: break;
    }
}

} // namespace pccl::cpu
