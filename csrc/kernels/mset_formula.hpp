// "mset64": an order-independent content hash, the third shared-state hash type next to simplehash and CRC32C.
//
// Why it exists: the reference hashes every shared-state tensor in a pass of its own before each synchronisation
// (ccoip_client_handler.cpp:383-416) and again after each transfer (:569-580).  simplehash's value depends on the
// order in which a fixed 960x256 launch geometry walks the buffer, so it cannot be produced by a kernel that visits the
// data in any other order.  mset64 is a sum (mod 2^64) of one well-mixed 64-bit value per 16-byte packet, keyed by
// the packet's index -- a multiset hash: any kernel that touches every packet exactly once, in whatever order and
// with whatever work partition, can accumulate it on the side.  The collective kernels do that while they write
// their output (all-reduce write-out, outer-step update), so the hash the next synchronisation needs is already
// known when the operation ends.
//
// Definition (all arithmetic mod 2^64, words little-endian, the last packet zero-padded to 16 bytes):
//   term(i, x, y, z, w) = mix( i*K0 + x*K1 + y*K2 + z*K3 + w*K4 ),   mix(m): m ^= m >> 29; m *= KM; m ^= m >> 32
//   raw   = sum of term(i, packet i) over all packets
//   mset64(data, n_bytes) = fmix64( raw ^ n_bytes*KL )
// Used only when every peer of a group advertises it (the master compares hash types); the default stays simplehash,
// which reference peers understand.
#pragma once
#include <cstdint>

#if defined(__CUDACC__)
#define PCCL_HD __host__ __device__ __forceinline__
#else
#define PCCL_HD inline
#endif

namespace pccl::mset {

constexpr uint64_t K0 = 0x9E3779B97F4A7C15ull, KM = 0xBF58476D1CE4E5B9ull, KL = 0xD6E8FEB86659FD93ull;
constexpr uint32_t K1 = 0x85EBCA77u, K2 = 0xC2B2AE3Du, K3 = 0x27D4EB2Fu, K4 = 0x165667B1u;

PCCL_HD uint64_t term(uint64_t index, uint32_t x, uint32_t y, uint32_t z, uint32_t w) {
    uint64_t m = index * K0 + static_cast<uint64_t>(x) * K1 + static_cast<uint64_t>(y) * K2 + static_cast<uint64_t>(z) * K3 +
                 static_cast<uint64_t>(w) * K4;
    m ^= m >> 29;
    m *= KM;
    m ^= m >> 32;
    return m;
}

PCCL_HD uint64_t finalize(uint64_t raw, uint64_t n_bytes) {
    uint64_t h = raw ^ (n_bytes * KL);
    h ^= h >> 33;
    h *= 0xff51afd7ed558ccdull;
    h ^= h >> 33;
    h *= 0xc4ceb9fe1a85ec53ull;
    h ^= h >> 33;
    return h;
}

} // namespace pccl::mset
