// C exports of the host kernels: they are the oracles the GPU tests compare against and the
// timed CPU baselines (reference CPU kernels: ccoip/src/cpp/reduce_kernels.cpp, quantize_kernels.cpp,
// simplehash/simplehash_cpu.cpp, crc32/crc32_cpu.cpp).
#include <cstring>

#include "../kernels/cpu/hash.hpp"
#include "../kernels/cpu/quantize.hpp"
#include "../kernels/cpu/reduce.hpp"
#include "pccl.h"

using namespace pccl;

namespace {
bool to_wire(int dtype, wire::DType &out) {
    static const wire::DType map[] = {wire::DType::U8,  wire::DType::I8,  wire::DType::I16, wire::DType::U16, wire::DType::U32,
                                      wire::DType::I32, wire::DType::U64, wire::DType::I64, wire::DType::F16, wire::DType::BF16,
                                      wire::DType::F32, wire::DType::F64, wire::DType::F8E4M3};
    if (dtype < 0 || dtype > 12) return false;
    out = map[dtype];
    return true;
}
bool to_op(int op, wire::RedOp &out) {
    switch (op) {
    case -1: out = wire::RedOp::Set; return true;
    case pcclSum: out = wire::RedOp::Sum; return true;
    case pcclAvg: out = wire::RedOp::Avg; return true;
    case pcclProd: out = wire::RedOp::Prod; return true;
    case pcclMax: out = wire::RedOp::Max; return true;
    case pcclMin: out = wire::RedOp::Min; return true;
    }
    return false;
}
} // namespace

extern "C" {

PCCL_EXPORT int pcclHostReduce(void *dst, const void *src, size_t count, int dtype, int op) {
    wire::DType t;
    wire::RedOp o;
    if (!to_wire(dtype, t) || !to_op(op, o)) return 1;
    cpu::reduce_inplace(dst, src, count, t, o);
    return 0;
}

PCCL_EXPORT int pcclHostScaleInv(void *dst, size_t count, int dtype, uint64_t divisor) {
    wire::DType t;
    if (!to_wire(dtype, t)) return 1;
    cpu::finalize_avg(dst, count, t, divisor);
    return 0;
}

// algorithm / code type as in pcclQuantizationOptions_t; value type fp32
PCCL_EXPORT size_t pcclHostQuantizedBytes(size_t count, int algorithm, int code_dtype) {
    cpu::QuantParams p;
    p.algo = static_cast<wire::QuantAlgo>(algorithm);
    if (!to_wire(code_dtype, p.code_dtype)) return 0;
    return cpu::quantized_size(p, count);
}

// meta_out: 2 doubles + 1 int64 = (min, max | scale, -) and zero point, for the per-chunk formats
PCCL_EXPORT int pcclHostQuantize(const float *src, void *out, size_t count, int algorithm, int code_dtype, double *meta_out) {
    cpu::QuantParams p;
    p.algo = static_cast<wire::QuantAlgo>(algorithm);
    p.value_dtype = wire::DType::F32;
    if (!to_wire(code_dtype, p.code_dtype)) return 1;
    if (!cpu::quant_supported(p.algo, p.value_dtype, p.code_dtype)) return 1;
    cpu::quant_analyze(src, count, p);
    cpu::quantize(src, out, count, p);
    if (meta_out) {
        meta_out[0] = p.algo == wire::QuantAlgo::ZeroPointScale ? p.scale : p.min;
        meta_out[1] = p.max;
        meta_out[2] = static_cast<double>(p.zero_point);
    }
    return 0;
}

PCCL_EXPORT int pcclHostDequantizeReduce(float *dst, const void *in, size_t count, int algorithm, int code_dtype, const double *meta,
                                         int op) {
    cpu::QuantParams p;
    p.algo = static_cast<wire::QuantAlgo>(algorithm);
    p.value_dtype = wire::DType::F32;
    wire::RedOp o;
    if (!to_wire(code_dtype, p.code_dtype) || !to_op(op, o)) return 1;
    if (meta) {
        if (p.algo == wire::QuantAlgo::ZeroPointScale) {
            p.scale = meta[0];
            p.zero_point = static_cast<int64_t>(meta[2]);
        } else {
            p.min = meta[0];
            p.max = meta[1];
        }
    }
    cpu::dequantize_reduce(dst, in, count, p, o);
    return 0;
}

PCCL_EXPORT uint32_t pcclHostCrc32cGeneric(const void *data, size_t n) { return cpu::crc32c_generic(data, n); }
PCCL_EXPORT void pcclHostCrc32cForceGeneric(int on) { cpu::crc32c_force_generic(on != 0); }
PCCL_EXPORT uint32_t pcclHostCrc32cCombine(uint32_t a, uint32_t b, uint64_t len_b) { return cpu::crc32c_combine(a, b, len_b); }

} // extern "C"
