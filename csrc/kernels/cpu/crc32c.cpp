#include <atomic>
#include <cstring>

#include "../crc_gf.hpp"
#include <algorithm>
#include <thread>
#include <vector>

#include "hash.hpp"

#if defined(__x86_64__)
#include <nmmintrin.h>
#endif

namespace pccl::cpu {

namespace {

constexpr uint32_t kPoly = gf::kCrc32cPoly; // reflected Castagnoli polynomial

struct Tables {
    uint32_t slice[8][256];
    // shift_long[k][b]: contribution of byte k of a CRC register after kLong more message bytes
    uint32_t shift_long[4][256];
    uint32_t shift_short[4][256];
};
constexpr size_t kLong = 8192, kShort = 256;

using gf::mul;
inline uint32_t gf_mul(uint32_t a, uint32_t b) { return gf::mul(a, b); }
inline uint32_t gf_x_pow_bits(uint64_t bits) { return gf::x_pow(bits); }

void build_shift_table(uint32_t table[4][256], size_t n_bytes) {
    const uint32_t factor = gf_x_pow_bits(static_cast<uint64_t>(n_bytes) * 8);
    for (int k = 0; k < 4; ++k)
        for (uint32_t b = 0; b < 256; ++b) table[k][b] = gf_mul(b << (8 * k), factor);
}

const Tables &tables() {
    static const Tables t = [] {
        Tables x{};
        for (uint32_t i = 0; i < 256; ++i) {
            uint32_t c = i;
            for (int k = 0; k < 8; ++k) c = (c >> 1) ^ ((c & 1u) ? kPoly : 0u);
            x.slice[0][i] = c;
        }
        for (uint32_t i = 0; i < 256; ++i)
            for (int k = 1; k < 8; ++k) x.slice[k][i] = (x.slice[k - 1][i] >> 8) ^ x.slice[0][x.slice[k - 1][i] & 0xffu];
        build_shift_table(x.shift_long, kLong);
        build_shift_table(x.shift_short, kShort);
        return x;
    }();
    return t;
}

inline uint32_t shift_by(const uint32_t table[4][256], uint32_t crc) {
    return table[0][crc & 0xff] ^ table[1][(crc >> 8) & 0xff] ^ table[2][(crc >> 16) & 0xff] ^ table[3][crc >> 24];
}

uint32_t raw_generic(uint32_t crc, const uint8_t *p, size_t n) {
    const Tables &t = tables();
    while (n && (reinterpret_cast<uintptr_t>(p) & 7)) {
        crc = (crc >> 8) ^ t.slice[0][(crc ^ *p++) & 0xff];
        --n;
    }
    while (n >= 8) {
        uint64_t w;
        std::memcpy(&w, p, 8);
        w ^= crc;
        crc = t.slice[7][w & 0xff] ^ t.slice[6][(w >> 8) & 0xff] ^ t.slice[5][(w >> 16) & 0xff] ^
              t.slice[4][(w >> 24) & 0xff] ^ t.slice[3][(w >> 32) & 0xff] ^ t.slice[2][(w >> 40) & 0xff] ^
              t.slice[1][(w >> 48) & 0xff] ^ t.slice[0][w >> 56];
        p += 8;
        n -= 8;
    }
    while (n--) crc = (crc >> 8) ^ t.slice[0][(crc ^ *p++) & 0xff];
    return crc;
}

#if defined(__x86_64__)
// Three independent crc32 instruction streams hide the 3-cycle latency; the partial results are
// merged with the precomputed "advance by N bytes" tables.
__attribute__((target("sse4.2"))) void three_way(uint64_t &c0, const uint8_t *&p, size_t &n, size_t block,
                                                 const uint32_t table[4][256]) {
    while (n >= 3 * block) {
        uint64_t c1 = 0, c2 = 0;
        const uint8_t *end = p + block;
        do {
            uint64_t a, b, c;
            std::memcpy(&a, p, 8);
            std::memcpy(&b, p + block, 8);
            std::memcpy(&c, p + 2 * block, 8);
            c0 = _mm_crc32_u64(c0, a);
            c1 = _mm_crc32_u64(c1, b);
            c2 = _mm_crc32_u64(c2, c);
            p += 8;
        } while (p < end);
        c0 = shift_by(table, static_cast<uint32_t>(c0)) ^ c1;
        c0 = shift_by(table, static_cast<uint32_t>(c0)) ^ c2;
        p += 2 * block;
        n -= 3 * block;
    }
}

__attribute__((target("sse4.2"))) uint32_t raw_sse42(uint32_t crc0, const uint8_t *p, size_t n) {
    const Tables &t = tables();
    uint64_t c0 = crc0;
    while (n && (reinterpret_cast<uintptr_t>(p) & 7)) {
        c0 = _mm_crc32_u8(static_cast<uint32_t>(c0), *p++);
        --n;
    }
    three_way(c0, p, n, kLong, t.shift_long);
    three_way(c0, p, n, kShort, t.shift_short);
    while (n >= 8) {
        uint64_t a;
        std::memcpy(&a, p, 8);
        c0 = _mm_crc32_u64(c0, a);
        p += 8;
        n -= 8;
    }
    while (n--) c0 = _mm_crc32_u8(static_cast<uint32_t>(c0), *p++);
    return static_cast<uint32_t>(c0);
}
#endif

std::atomic<bool> g_force_generic{false};

} // namespace

bool crc32c_hardware_available() {
#if defined(__x86_64__)
    return __builtin_cpu_supports("sse4.2");
#else
    return false;
#endif
}

void crc32c_force_generic(bool on) { g_force_generic = on; }

uint32_t crc32c_generic(const void *data, size_t n_bytes) {
    return ~raw_generic(0xFFFFFFFFu, static_cast<const uint8_t *>(data), n_bytes);
}

namespace {
uint32_t crc32c_one_thread(const void *data, size_t n_bytes) {
#if defined(__x86_64__)
    if (!g_force_generic && crc32c_hardware_available())
        return ~raw_sse42(0xFFFFFFFFu, static_cast<const uint8_t *>(data), n_bytes);
#endif
    return crc32c_generic(data, n_bytes);
}
} // namespace

// Large buffers are cut into one piece per thread; crc(A||B) follows from crc(A), crc(B) and the length of B (the same
// algebra the device kernel uses to combine its lanes), so the pieces are hashed independently.
uint32_t crc32c(const void *data, size_t n_bytes) {
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    size_t n_threads = std::min<size_t>(std::min(8u, std::max(1u, hw / 2)), n_bytes / (4u << 20));
    if (hash_thread_cap()) n_threads = std::min<size_t>(n_threads, hash_thread_cap());
    if (n_threads <= 1) return crc32c_one_thread(data, n_bytes);
    const auto *bytes = static_cast<const uint8_t *>(data);
    const size_t per = ((n_bytes + n_threads - 1) / n_threads + 63) / 64 * 64;
    std::vector<uint32_t> part(n_threads, 0);
    std::vector<size_t> len(n_threads, 0);
    std::vector<std::thread> pool;
    for (size_t t = 0; t < n_threads; ++t) {
        const size_t lo = std::min(n_bytes, t * per), hi = std::min(n_bytes, lo + per);
        len[t] = hi - lo;
        pool.emplace_back([&, t, lo] { part[t] = crc32c_one_thread(bytes + lo, len[t]); });
    }
    for (auto &th : pool) th.join();
    uint32_t crc = part[0];
    for (size_t t = 1; t < n_threads; ++t)
        if (len[t]) crc = crc32c_combine(crc, part[t], len[t]);
    return crc;
}

uint32_t crc32c_combine(uint32_t crc_a, uint32_t crc_b, uint64_t len_b) {
    // crc(A||B) = crc(A) * x^(8 len_b) + crc(B) in the affine sense: the init/xor-out masks of the
    // two parts cancel except for one, which the algebra below accounts for.
    return gf_mul(crc_a, gf_x_pow_bits(len_b * 8)) ^ crc_b;
}

} // namespace pccl::cpu
