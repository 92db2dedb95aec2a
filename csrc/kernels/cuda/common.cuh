// Shared device helpers: element-type traits, 16-byte vector access, system-scope flag access.
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>

#include <cstdint>

#include "pccl.h"

namespace pccl::cuda {

enum : int { OP_SUM = 0, OP_PROD = 1, OP_MAX = 2, OP_MIN = 3 };

// Accumulator / arithmetic traits per element type.
template <typename T> struct Elem {
    using acc_t = T;
    static __device__ __forceinline__ acc_t load(T v) { return v; }
    static __device__ __forceinline__ T store(acc_t v) { return v; }
};
template <> struct Elem<__half> {
    using acc_t = float;
    static __device__ __forceinline__ float load(__half v) { return __half2float(v); }
    static __device__ __forceinline__ __half store(float v) { return __float2half_rn(v); }
};
template <> struct Elem<__nv_bfloat16> {
    using acc_t = float;
    static __device__ __forceinline__ float load(__nv_bfloat16 v) { return __bfloat162float(v); }
    static __device__ __forceinline__ __nv_bfloat16 store(float v) { return __float2bfloat16_rn(v); }
};

template <int OP, typename A> __device__ __forceinline__ A combine(A a, A b) {
    if constexpr (OP == OP_SUM) return a + b;
    if constexpr (OP == OP_PROD) return a * b;
    if constexpr (OP == OP_MAX) return a > b ? a : b;
    return a < b ? a : b;
}

// 16-byte packet of elements
template <typename T> struct alignas(16) Pack {
    static constexpr int N = 16 / sizeof(T);
    T v[N];
};

__device__ __forceinline__ uint4 ld_global_16(const void *p) { return *reinterpret_cast<const uint4 *>(p); }
__device__ __forceinline__ void st_global_16(void *p, uint4 v) { *reinterpret_cast<uint4 *>(p) = v; }

// Streaming accesses: data touched once, keep it out of L1.
__device__ __forceinline__ uint4 ld_stream_16(const void *p) {
    uint4 v;
    asm volatile("ld.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void st_stream_16(void *p, uint4 v) {
    asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// Peer memory written by another GPU during this kernel: must not be served from a stale L1 line.
__device__ __forceinline__ uint4 ld_peer_16(const void *p) {
    uint4 v;
    asm volatile("ld.relaxed.sys.global.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                 : "l"(p)
                 : "memory");
    return v;
}
__device__ __forceinline__ uint32_t ld_peer_4(const void *p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.L1::no_allocate.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint2 ld_peer_8(const void *p) {
    uint2 v;
    asm volatile("ld.relaxed.sys.global.L1::no_allocate.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p) : "memory");
    return v;
}

__device__ __forceinline__ void st_flag_release(uint64_t *p, uint64_t v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_flag_acquire(const uint64_t *p) {
    uint64_t v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

// host-side dispatch helpers ---------------------------------------------------------------
template <typename F> inline bool dispatch_dtype(int dtype, F &&f) {
    switch (dtype) {
    case pcclUint8: f(uint8_t{}); return true;
    case pcclInt8: f(int8_t{}); return true;
    case pcclInt16: f(int16_t{}); return true;
    case pcclUint16: f(uint16_t{}); return true;
    case pcclUint32: f(uint32_t{}); return true;
    case pcclInt32: f(int32_t{}); return true;
    case pcclUint64: f(uint64_t{}); return true;
    case pcclInt64: f(int64_t{}); return true;
    case pcclFloat16: f(__half{}); return true;
    case pcclBFloat16: f(__nv_bfloat16{}); return true;
    case pcclFloat: f(float{}); return true;
    case pcclDouble: f(double{}); return true;
    default: return false;
    }
}
// pcclRedOp_t -> kernel operator; avg is sum followed by a division
inline int kernel_op(int op, bool &is_avg) {
    is_avg = op == pcclAvg;
    switch (op) {
    case pcclSum:
    case pcclAvg: return OP_SUM;
    case pcclProd: return OP_PROD;
    case pcclMax: return OP_MAX;
    case pcclMin: return OP_MIN;
    default: return -1;
    }
}
template <typename F> inline bool dispatch_op(int kop, F &&f) {
    switch (kop) {
    case OP_SUM: f(std::integral_constant<int, OP_SUM>{}); return true;
    case OP_PROD: f(std::integral_constant<int, OP_PROD>{}); return true;
    case OP_MAX: f(std::integral_constant<int, OP_MAX>{}); return true;
    case OP_MIN: f(std::integral_constant<int, OP_MIN>{}); return true;
    default: return false;
    }
}

} // namespace pccl::cuda
