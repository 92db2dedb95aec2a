// Vectorised float32 / 8-bit-code paths of the host quantisers; same results as the generic templates in quantize.cpp,
// bit for bit (see quantize_fast.cpp).
#pragma once
#include <cstddef>
#include <cstdint>

#include "quantize.hpp"

namespace pccl::cpu::fast {

void extrema_f32(const float *x, size_t n, float &lo, float &hi);
void minmax_quant_u8(const float *x, uint8_t *q, size_t n, float mn, float mx);
void minmax_dequant_u8_set(float *dst, const uint8_t *q, size_t n, float mn, float mx);
void minmax_dequant_u8_add(float *dst, const uint8_t *q, size_t n, float mn, float mx);
void zps_quant_u8(const float *x, uint8_t *q, size_t n, double scale, int64_t zp);
void zps_dequant_u8_set(float *dst, const uint8_t *q, size_t n, double scale, int64_t zp);
void zps_dequant_u8_add(float *dst, const uint8_t *q, size_t n, double scale, int64_t zp);
void block_minmax_quant(const float *x, uint8_t *out, size_t n);
void block_minmax_dequant_set_fma(float *dst, const uint8_t *in, size_t n); // only when cpu_has_fma()
void block_minmax_dequant_add_fma(float *dst, const uint8_t *in, size_t n);
void block_fp8_quant(const float *x, uint8_t *out, size_t n);
void block_fp8_dequant_set(float *dst, const uint8_t *in, size_t n);
void block_fp8_dequant_add(float *dst, const uint8_t *in, size_t n);
bool cpu_has_fma();

// the conversions used inside the two fp8 loops, exported for the exhaustive comparison in the unit tests
uint8_t f32_to_fp8e4m3_branchless(float f);
float fp8e4m3_to_f32_branchless(uint8_t c);

} // namespace pccl::cpu::fast
