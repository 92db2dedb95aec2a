// Software conversions for the 16-bit float formats on the host.
#pragma once
#include <bit>
#include <cstdint>

namespace pccl::cpu {

inline float bf16_to_f32(uint16_t h) { return std::bit_cast<float>(static_cast<uint32_t>(h) << 16); }
inline uint16_t f32_to_bf16(float f) { // round to nearest even, NaN stays NaN
    uint32_t u = std::bit_cast<uint32_t>(f);
    if ((u & 0x7fffffffu) > 0x7f800000u) return static_cast<uint16_t>((u >> 16) | 0x40);
    u += 0x7fffu + ((u >> 16) & 1u);
    return static_cast<uint16_t>(u >> 16);
}

inline float f16_to_f32(uint16_t h) {
    const uint32_t sign = static_cast<uint32_t>(h & 0x8000u) << 16;
    uint32_t exp = (h >> 10) & 0x1fu, man = h & 0x3ffu;
    uint32_t out;
    if (exp == 0) {
        if (man == 0) {
            out = sign;
        } else { // subnormal: normalise
            int shift = 0;
            while (!(man & 0x400u)) {
                man <<= 1;
                ++shift;
            }
            man &= 0x3ffu;
            out = sign | ((127 - 15 - shift + 1) << 23) | (man << 13);
        }
    } else if (exp == 31) {
        out = sign | 0x7f800000u | (man << 13);
    } else {
        out = sign | ((exp + 127 - 15) << 23) | (man << 13);
    }
    return std::bit_cast<float>(out);
}

inline uint16_t f32_to_f16(float f) { // round to nearest even
    const uint32_t u = std::bit_cast<uint32_t>(f);
    const uint16_t sign = static_cast<uint16_t>((u >> 16) & 0x8000u);
    const uint32_t abs = u & 0x7fffffffu;
    if (abs > 0x7f800000u) return static_cast<uint16_t>(sign | 0x7e00u);           // NaN
    if (abs >= 0x477ff000u) return static_cast<uint16_t>(sign | 0x7c00u);          // overflow -> inf (>= 65520)
    if (abs < 0x33000001u) return sign;                                            // underflow -> 0
    int exp = static_cast<int>(abs >> 23) - 127 + 15;
    uint32_t man = (abs & 0x7fffffu) | 0x800000u;
    if (exp <= 0) { // subnormal result
        const int shift = 14 - exp; // 13 + (1 - exp)
        const uint32_t rounded = man >> shift;
        const uint32_t rem = man & ((1u << shift) - 1), half = 1u << (shift - 1);
        uint32_t r = rounded + ((rem > half || (rem == half && (rounded & 1u))) ? 1u : 0u);
        return static_cast<uint16_t>(sign | r);
    }
    uint32_t r = (static_cast<uint32_t>(exp) << 10) | ((man >> 13) & 0x3ffu);
    const uint32_t rem = man & 0x1fffu;
    if (rem > 0x1000u || (rem == 0x1000u && (r & 1u))) ++r; // may carry into exponent, which is correct
    return static_cast<uint16_t>(sign | r);
}

} // namespace pccl::cpu
