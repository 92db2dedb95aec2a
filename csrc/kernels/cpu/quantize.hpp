// Host-side quantisation for the TCP data plane, and the oracle for the GPU quantisers.
//
//  MinMax          reference formulas (ccoip/src/cpp/quantize_kernels.cpp:41-58, quantize_kernels.hpp:15-22):
//                  one (min,max) per chunk, q = trunc((x-min)/(max-min) * QMAX) in double,
//                  x' = min + T(q / QMAX) * (max-min).  (The reference seeds the max scan with the
//                  smallest positive value; here it is seeded with the lowest value.)
//  ZeroPointScale  native replacement of the missing piquant library (call sites in the reference:
//                  ccoip/src/cpp/quantize.cpp:22-57, reduce_kernels.cpp:370-405): scale = (max-min)/(QMAX-QMIN),
//                  zp = round(QMIN - min/scale), q = clamp(round(x/scale) + zp), x' = (q - zp) * scale.
//                  A constant chunk is encoded as scale = value, q = 1 so it survives exactly.
//  BlockMinMax     (extension) uint8 code words, one fp32 (min,max) pair per 128 elements.
//  BlockFp8        (extension) fp8-e4m3 code words, one fp32 scale per 128 elements.
//
// Block formats are self-describing: payload = [per-block metadata][code words].
#pragma once
#include <cstddef>
#include <cstdint>
#include <vector>

#include "../../wire/packets.hpp"

namespace pccl::cpu {

constexpr size_t kQuantBlock = 128;

struct QuantParams {
    wire::QuantAlgo algo = wire::QuantAlgo::None;
    wire::DType value_dtype = wire::DType::F32; // F32 or F64
    wire::DType code_dtype = wire::DType::U8;
    double min = 0, max = 0;                    // MinMax
    double scale = 1;                           // ZeroPointScale
    int64_t zero_point = 0;
};

bool quant_supported(wire::QuantAlgo algo, wire::DType value_dtype, wire::DType code_dtype);
// bytes occupied on the wire by `count` quantised elements
size_t quantized_size(const QuantParams &p, size_t count);
// scans the chunk and fills min/max or scale/zero point
void quant_analyze(const void *src, size_t count, QuantParams &p);
void quantize(const void *src, void *dst, size_t count, const QuantParams &p);
// dst[i] = dst[i] (op) dequant(q[i]);  op == Set stores the dequantised value
void dequantize_reduce(void *dst, const void *q, size_t count, const QuantParams &p, wire::RedOp op);

wire::P2PDequantMeta to_wire(const QuantParams &p, uint64_t tag);
bool from_wire(const wire::P2PDequantMeta &m, QuantParams &p); // fills the numeric fields only

// Tests only: true routes everything through the generic (scalar, branchy) templates, which define the arithmetic; the
// vectorised float32 / 8-bit paths must reproduce them bit for bit.  Also settable with PCCL_QUANT_GENERIC=1.
void quant_force_generic(bool on);

float fp8e4m3_to_f32(uint8_t v);
uint8_t f32_to_fp8e4m3(float f); // round to nearest even, saturating at +-448

} // namespace pccl::cpu
