// Element-wise device kernels used when device buffers travel over the TCP data plane and by
// the tests: dst (op)= src, AVG finalisation, block-scaled quantise / dequantise-reduce.
// Host oracles: csrc/kernels/cpu/reduce.cpp, csrc/kernels/cpu/quantize.cpp.  Reference CPU
// versions: ccoip/src/cpp/reduce_kernels.cpp:56-215,437-510 and quantize_kernels.cpp:24-217.
#include "common.cuh"
#include "pccl_kernels.h"
#include "quant.cuh"

namespace pccl::cuda {

constexpr int kEwThreads = 256;

inline int ew_grid(uint64_t work_items) {
    const uint64_t blocks = (work_items + kEwThreads - 1) / kEwThreads;
    return static_cast<int>(std::max<uint64_t>(1, std::min<uint64_t>(blocks, 148ull * 8)));
}

template <typename T, int OP> __global__ void __launch_bounds__(kEwThreads) reduce_kernel(T *__restrict__ dst, const T *__restrict__ src, uint64_t count) {
    constexpr int N = 16 / sizeof(T);
    const bool aligned = ((reinterpret_cast<uintptr_t>(dst) | reinterpret_cast<uintptr_t>(src)) & 15) == 0;
    const uint64_t n_pack = aligned ? count / N : 0;
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x, tid = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    for (uint64_t v = tid; v < n_pack; v += stride) {
        const uint4 a = ld_stream_16(dst + v * N), b = ld_stream_16(src + v * N);
        const Pack<T> pa = *reinterpret_cast<const Pack<T> *>(&a), pb = *reinterpret_cast<const Pack<T> *>(&b);
        Pack<T> out;
#pragma unroll
        for (int i = 0; i < N; ++i) out.v[i] = Elem<T>::store(combine<OP>(Elem<T>::load(pa.v[i]), Elem<T>::load(pb.v[i])));
        st_stream_16(dst + v * N, *reinterpret_cast<uint4 *>(&out));
    }
    for (uint64_t e = n_pack * N + tid; e < count; e += stride)
        dst[e] = Elem<T>::store(combine<OP>(Elem<T>::load(dst[e]), Elem<T>::load(src[e])));
}

template <typename T> __global__ void __launch_bounds__(kEwThreads) scale_inv_kernel(T *__restrict__ dst, uint64_t count, uint64_t divisor) {
    using A = typename Elem<T>::acc_t;
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    for (uint64_t e = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; e < count; e += stride)
        dst[e] = Elem<T>::store(Elem<T>::load(dst[e]) / static_cast<A>(divisor));
}

// out = [block records][code bytes]; one warp per 128-element block
template <int ALGO> __global__ void __launch_bounds__(kEwThreads) quantize_kernel(const float *__restrict__ src, uint8_t *__restrict__ out, uint64_t count) {
    constexpr int MB = QuantMeta<ALGO>::kBytes;
    const uint64_t n_blocks = (count + PCCL_K_QUANT_BLOCK - 1) / PCCL_K_QUANT_BLOCK;
    uint8_t *codes = out + n_blocks * MB;
    const int lane = threadIdx.x & 31;
    const uint64_t warps = uint64_t(gridDim.x) * (blockDim.x / 32), warp = uint64_t(blockIdx.x) * (blockDim.x / 32) + threadIdx.x / 32;
    const bool aligned = (reinterpret_cast<uintptr_t>(src) & 15) == 0;
    for (uint64_t b = warp; b < n_blocks; b += warps) {
        const uint64_t e = b * PCCL_K_QUANT_BLOCK + lane * 4;
        float x[4] = {0, 0, 0, 0};
        int valid = 0;
        if (e < count) {
            valid = count - e >= 4 ? 4 : static_cast<int>(count - e);
            if (valid == 4 && aligned) {
                const uint4 raw = ld_stream_16(src + e);
                x[0] = __uint_as_float(raw.x), x[1] = __uint_as_float(raw.y), x[2] = __uint_as_float(raw.z), x[3] = __uint_as_float(raw.w);
            } else {
                for (int i = 0; i < valid; ++i) x[i] = src[e + i];
            }
        }
        const uint32_t packed = quantize4<ALGO>(x, valid, out + b * MB, lane);
        for (int i = 0; i < valid; ++i) codes[e + i] = static_cast<uint8_t>(packed >> (8 * i));
    }
}

// OP < 0: store
template <int ALGO, int OP> __global__ void __launch_bounds__(kEwThreads) dequantize_reduce_kernel(float *__restrict__ dst, const uint8_t *__restrict__ in, uint64_t count) {
    constexpr int MB = QuantMeta<ALGO>::kBytes;
    const uint64_t n_blocks = (count + PCCL_K_QUANT_BLOCK - 1) / PCCL_K_QUANT_BLOCK;
    const uint8_t *codes = in + n_blocks * MB;
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    for (uint64_t e = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; e < count; e += stride) {
        const uint64_t b = e / PCCL_K_QUANT_BLOCK;
        uint32_t mlo, mhi = 0;
        if constexpr (MB == 8) {
            const uint2 m = *reinterpret_cast<const uint2 *>(in + b * MB);
            mlo = m.x, mhi = m.y;
        } else {
            mlo = *reinterpret_cast<const uint32_t *>(in + b * MB);
        }
        float y[4];
        dequantize4<ALGO>(codes[e], mlo, mhi, y);
        if constexpr (OP < 0)
            dst[e] = y[0];
        else
            dst[e] = combine<OP>(dst[e], y[0]);
    }
}

} // namespace pccl::cuda

using namespace pccl::cuda;

extern "C" int pcclKernelReduce(void *dst, const void *src, size_t count, int dtype, int op, void *stream_v) {
    if (count == 0) return 0;
    bool is_avg = false;
    const int kop = kernel_op(op, is_avg);
    if (kop < 0) return cudaErrorInvalidValue;
    const auto stream = static_cast<cudaStream_t>(stream_v);
    bool launched = false;
    dispatch_dtype(dtype, [&](auto tag) {
        using T = decltype(tag);
        dispatch_op(kop, [&](auto op_tag) {
            constexpr int OP = decltype(op_tag)::value;
            reduce_kernel<T, OP><<<ew_grid(count * sizeof(T) / 16 + 1), kEwThreads, 0, stream>>>(static_cast<T *>(dst), static_cast<const T *>(src), count);
            launched = true;
        });
    });
    return launched ? cudaGetLastError() : cudaErrorInvalidValue;
}

extern "C" int pcclKernelScaleInv(void *dst, size_t count, int dtype, uint64_t divisor, void *stream_v) {
    if (count == 0 || divisor <= 1) return 0;
    const auto stream = static_cast<cudaStream_t>(stream_v);
    bool launched = false;
    dispatch_dtype(dtype, [&](auto tag) {
        using T = decltype(tag);
        scale_inv_kernel<T><<<ew_grid(count), kEwThreads, 0, stream>>>(static_cast<T *>(dst), count, divisor);
        launched = true;
    });
    return launched ? cudaGetLastError() : cudaErrorInvalidValue;
}

extern "C" size_t pcclKernelQuantizedBytes(size_t count, int algorithm) {
    const size_t n_blocks = (count + PCCL_K_QUANT_BLOCK - 1) / PCCL_K_QUANT_BLOCK;
    if (algorithm == QUANT_BLOCK_MINMAX) return n_blocks * 8 + count;
    if (algorithm == QUANT_BLOCK_FP8) return n_blocks * 4 + count;
    return 0;
}

extern "C" int pcclKernelQuantize(const float *src, void *out, size_t count, int algorithm, void *stream_v) {
    if (count == 0) return 0;
    const auto stream = static_cast<cudaStream_t>(stream_v);
    const uint64_t n_blocks = (count + PCCL_K_QUANT_BLOCK - 1) / PCCL_K_QUANT_BLOCK;
    const int grid = ew_grid(n_blocks * 32);
    if (algorithm == QUANT_BLOCK_MINMAX)
        quantize_kernel<QUANT_BLOCK_MINMAX><<<grid, kEwThreads, 0, stream>>>(src, static_cast<uint8_t *>(out), count);
    else if (algorithm == QUANT_BLOCK_FP8)
        quantize_kernel<QUANT_BLOCK_FP8><<<grid, kEwThreads, 0, stream>>>(src, static_cast<uint8_t *>(out), count);
    else
        return cudaErrorInvalidValue;
    return cudaGetLastError();
}

extern "C" int pcclKernelDequantizeReduce(float *dst, const void *in, size_t count, int algorithm, int op, void *stream_v) {
    if (count == 0) return 0;
    const auto stream = static_cast<cudaStream_t>(stream_v);
    const int grid = ew_grid(count);
    const auto *src = static_cast<const uint8_t *>(in);
    bool is_avg = false;
    const int kop = op < 0 ? -1 : kernel_op(op, is_avg);
    if (op >= 0 && kop < 0) return cudaErrorInvalidValue;
    auto launch = [&](auto algo_tag) {
        constexpr int ALGO = decltype(algo_tag)::value;
        if (kop < 0) {
            dequantize_reduce_kernel<ALGO, -1><<<grid, kEwThreads, 0, stream>>>(dst, src, count);
            return;
        }
        dispatch_op(kop, [&](auto op_tag) {
            dequantize_reduce_kernel<ALGO, decltype(op_tag)::value><<<grid, kEwThreads, 0, stream>>>(dst, src, count);
        });
    };
    if (algorithm == QUANT_BLOCK_MINMAX)
        launch(std::integral_constant<int, QUANT_BLOCK_MINMAX>{});
    else if (algorithm == QUANT_BLOCK_FP8)
        launch(std::integral_constant<int, QUANT_BLOCK_FP8>{});
    else
        return cudaErrorInvalidValue;
    return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// DiLoCo outer step (reference does this with torch ops on the CPU:
// python/examples/nanogpt_diloco/sync_diloco.py:480-530): pseudo-gradient and the outer Nesterov-SGD
// update, each one pass over memory.
// ---------------------------------------------------------------------------------------------
namespace pccl::cuda {

__global__ void __launch_bounds__(kEwThreads) pseudo_grad_kernel(float *__restrict__ out, const float *__restrict__ outer,
                                                                  const float *__restrict__ inner, uint64_t count) {
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x, tid = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x;
    const bool aligned = ((reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(outer) | reinterpret_cast<uintptr_t>(inner)) & 15) == 0;
    const uint64_t n4 = aligned ? count / 4 : 0;
    for (uint64_t v = tid; v < n4; v += stride) {
        const uint4 a = ld_stream_16(outer + v * 4), b = ld_stream_16(inner + v * 4);
        uint4 r;
        r.x = __float_as_uint(__uint_as_float(a.x) - __uint_as_float(b.x));
        r.y = __float_as_uint(__uint_as_float(a.y) - __uint_as_float(b.y));
        r.z = __float_as_uint(__uint_as_float(a.z) - __uint_as_float(b.z));
        r.w = __float_as_uint(__uint_as_float(a.w) - __uint_as_float(b.w));
        st_stream_16(out + v * 4, r);
    }
    for (uint64_t e = n4 * 4 + tid; e < count; e += stride) out[e] = outer[e] - inner[e];
}

// torch.optim.SGD(momentum=mu, nesterov=True) semantics: buf = first ? g : mu*buf + g;
// p -= lr * (g + mu*buf); the new parameters are also written to `inner` (next inner phase starts there)
__global__ void __launch_bounds__(kEwThreads) outer_nesterov_kernel(float *__restrict__ params, float *__restrict__ buf,
                                                                     const float *__restrict__ grad, float *__restrict__ inner,
                                                                     uint64_t count, float lr, float mu, int first) {
    const uint64_t stride = uint64_t(gridDim.x) * blockDim.x;
    for (uint64_t e = uint64_t(blockIdx.x) * blockDim.x + threadIdx.x; e < count; e += stride) {
        const float g = grad[e];
        const float b = first ? g : fmaf(mu, buf[e], g);
        const float p = params[e] - lr * fmaf(mu, b, g);
        buf[e] = b;
        params[e] = p;
        if (inner) inner[e] = p;
    }
}

} // namespace pccl::cuda

extern "C" PCCL_EXPORT int pcclKernelPseudoGrad(float *out, const float *outer, const float *inner, size_t count, void *stream_v) {
    if (count == 0) return 0;
    pccl::cuda::pseudo_grad_kernel<<<ew_grid(count / 4 + 1), kEwThreads, 0, static_cast<cudaStream_t>(stream_v)>>>(out, outer, inner, count);
    return cudaGetLastError();
}

extern "C" PCCL_EXPORT int pcclKernelOuterNesterov(float *params, float *momentum, const float *grad, float *inner, size_t count, float lr,
                                                   float mu, int first, void *stream_v) {
    if (count == 0) return 0;
    pccl::cuda::outer_nesterov_kernel<<<ew_grid(count), kEwThreads, 0, static_cast<cudaStream_t>(stream_v)>>>(params, momentum, grad, inner, count, lr,
                                                                                                          mu, first);
    return cudaGetLastError();
}
