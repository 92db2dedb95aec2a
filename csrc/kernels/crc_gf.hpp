// GF(2) polynomial arithmetic modulo the (reflected) Castagnoli polynomial, shared by the host
// CRC32C, the device CRC32C and their tests.  Bit 31 of a word is the coefficient of x^0.
#pragma once
#include <cstdint>

#ifdef __CUDACC__
#define PCCL_HD __host__ __device__ __forceinline__
#else
#define PCCL_HD inline
#endif

namespace pccl::gf {

constexpr uint32_t kCrc32cPoly = 0x82F63B78u;

PCCL_HD uint32_t mul(uint32_t a, uint32_t b) { // a * b mod P
    uint32_t p = 0;
    for (int i = 0; i < 32; ++i) {
        if (a & 0x80000000u) p ^= b;
        a <<= 1;
        b = (b >> 1) ^ ((b & 1u) ? kCrc32cPoly : 0u);
    }
    return p;
}

PCCL_HD uint32_t x_pow(uint64_t bits) { // x^bits mod P
    uint32_t result = 0x80000000u, base = 0x40000000u;
    while (bits) {
        if (bits & 1) result = mul(result, base);
        base = mul(base, base);
        bits >>= 1;
    }
    return result;
}

} // namespace pccl::gf
