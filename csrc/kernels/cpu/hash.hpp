// Content hashes of the shared-state protocol on the host.
//
// simplehash: the value is defined by the reference's CUDA kernel pair (ccoip/src/cuda/simplehash_cuda.cu:95-267)
// and its CPU twin (ccoip/src/cpp/simplehash/simplehash_cpu.cpp:75-200): a non-associative combine
// applied in the order fixed by a virtual launch of min(ceil(n_vec/256), 960) blocks x 256 threads, then
// warp-shuffle trees, then a 256-thread final pass, then tail words and tail bytes.  This file
// re-derives that order with plain loops; golden values in tests/cpp/test_hash.cpp.
//
// crc32c: Castagnoli CRC (reflected polynomial 0x82F63B78), init and final xor 0xFFFFFFFF.
#pragma once
#include <cstddef>
#include <cstdint>

namespace pccl::cpu {

constexpr uint32_t kSimpleHashMaxBlocks = 960;
constexpr uint32_t kSimpleHashThreads = 256;

inline uint32_t hash_combine(uint32_t a, uint32_t b) {
    a ^= b + 0x9e3779b1u;
    a = (a << 7) | (a >> 25);
    a *= 0x85ebca6bu;
    return a;
}

// Upper bound on the threads one hash call of THIS thread may use (0 = no bound).  A caller that hashes many tensors at once,
// one per thread, sets 1 so that the calls do not each start a crew of their own.
void hash_thread_cap(unsigned max_threads);
unsigned hash_thread_cap();

// Hashes `n` host buffers with the given hash type (0 simplehash, 1 CRC32C, 2 mset64), several buffers at a time when there
// are several: results in out[i].
void hash_many(const void *const *data, const size_t *n_bytes, size_t n, int hash_type, uint64_t *out);

uint64_t simplehash(const void *data, size_t n_bytes);

// Folds 256 per-thread values the way the device block reduction does (8 warp trees, then one
// tree over the warp results padded with zeros).  Exposed for the GPU tests.
uint32_t simplehash_block_fold(const uint32_t *thread_values);

// mset64: order-independent sum of per-packet terms (csrc/kernels/mset_formula.hpp), the hash the collective kernels can
// produce while they write their output.  mset64_raw: the unfinalised sum of a span whose first packet has the given index.
uint64_t mset64(const void *data, size_t n_bytes);
uint64_t mset64_raw(const void *data, size_t n_bytes, uint64_t first_packet);

uint32_t crc32c(const void *data, size_t n_bytes);
uint32_t crc32c_generic(const void *data, size_t n_bytes); // table driven, no special instructions
// CRC of the concatenation A||B from crc(A), crc(B) and len(B)
uint32_t crc32c_combine(uint32_t crc_a, uint32_t crc_b, uint64_t len_b);
// test hook: true forces the table-driven path regardless of CPU features
void crc32c_force_generic(bool on);
bool crc32c_hardware_available();

} // namespace pccl::cpu
