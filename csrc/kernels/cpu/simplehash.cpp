#include <algorithm>
#include <atomic>
#include <cstring>
#include <thread>
#include <vector>

#include "hash.hpp"

namespace pccl::cpu {

namespace {
thread_local unsigned t_thread_cap = 0;
} // namespace

void hash_thread_cap(unsigned max_threads) { t_thread_cap = max_threads; }
unsigned hash_thread_cap() { return t_thread_cap; }

void hash_many(const void *const *data, const size_t *n_bytes, size_t n, int hash_type, uint64_t *out) {
    auto one = [&](size_t i) {
        out[i] = hash_type == 0 ? simplehash(data[i], n_bytes[i]) : hash_type == 1 ? crc32c(data[i], n_bytes[i]) : mset64(data[i], n_bytes[i]);
    };
    size_t total = 0;
    for (size_t i = 0; i < n; ++i) total += n_bytes[i];
    const unsigned crew = std::min<unsigned>({8u, std::max(1u, std::thread::hardware_concurrency() / 2), static_cast<unsigned>(n)});
    if (n < 2 || crew < 2 || total < (8u << 20)) { // one buffer, or little data: the calls thread themselves where it pays
        for (size_t i = 0; i < n; ++i) one(i);
        return;
    }
    std::atomic<size_t> next{0};
    std::vector<std::thread> pool;
    for (unsigned t = 0; t < crew; ++t)
        pool.emplace_back([&] {
            hash_thread_cap(1);
            for (size_t i = next++; i < n; i = next++) one(i);
        });
    for (auto &th : pool) th.join();
}

namespace {

// One shuffle tree over 32 lanes: at every level lane i takes lane i+offset (lanes without a partner combine with
// themselves -- that is what __shfl_down_sync hands back), and only lane 0 is read afterwards.  Lane 0 at the end depends on
// lanes 0..offset-1 of each level only, and those always have a partner, so the tree is computed on a shrinking prefix.
uint32_t warp_fold(const uint32_t *lanes) {
    uint32_t v[32];
    std::memcpy(v, lanes, sizeof v);
    for (int offset = 16; offset > 0; offset >>= 1)
        for (int i = 0; i < offset; ++i) v[i] = hash_combine(v[i], v[i + offset]);
    return v[0];
}

// The slab of one virtual block: virtual thread t folds packets t, t+256, t+512, ...  Walking the slab row by row (256
// consecutive packets, one per virtual thread) keeps 256 independent chains going, which the compiler turns into vector
// code (interleaved loads, packed multiply and rotate): several times the speed of following one chain at a time, same value.
#if defined(__x86_64__) && defined(__GNUC__) && !defined(__clang__) && !defined(__SANITIZE_ADDRESS__) && !defined(__SANITIZE_THREAD__)
__attribute__((target_clones("arch=x86-64-v4", "arch=x86-64-v3", "default")))
#endif
uint32_t block_partial(const uint32_t *words, uint64_t vec_begin, uint64_t vec_end) {
    uint32_t h[kSimpleHashThreads];
    for (uint32_t t = 0; t < kSimpleHashThreads; ++t) h[t] = 0;
    uint64_t i = vec_begin;
    for (; i + kSimpleHashThreads <= vec_end; i += kSimpleHashThreads) {
        const uint32_t *w = words + i * 4;
        for (uint32_t t = 0; t < kSimpleHashThreads; ++t) {
            uint32_t x = h[t];
            x = hash_combine(x, w[4 * t]);
            x = hash_combine(x, w[4 * t + 1]);
            x = hash_combine(x, w[4 * t + 2]);
            x = hash_combine(x, w[4 * t + 3]);
            h[t] = x;
        }
    }
    const uint32_t *w = words + i * 4;
    for (uint32_t t = 0; i + t < vec_end; ++t) {
        uint32_t x = h[t];
        x = hash_combine(x, w[4 * t]);
        x = hash_combine(x, w[4 * t + 1]);
        x = hash_combine(x, w[4 * t + 2]);
        x = hash_combine(x, w[4 * t + 3]);
        h[t] = x;
    }
    return simplehash_block_fold(h);
}

} // namespace

uint32_t simplehash_block_fold(const uint32_t *thread_values) {
    uint32_t warp_results[32] = {0};
    for (int w = 0; w < static_cast<int>(kSimpleHashThreads / 32); ++w) warp_results[w] = warp_fold(thread_values + w * 32);
    return warp_fold(warp_results);
}

uint64_t simplehash(const void *data, size_t n_bytes) {
    if (n_bytes == 0) return 0;
    const auto *bytes = static_cast<const uint8_t *>(data);
    const size_t n_words = n_bytes / 4, n_vec = n_words / 4;
    uint32_t result = 0;

    // unaligned input is legal on the host: go through a word view only when it is aligned
    std::vector<uint32_t> aligned_copy;
    const uint32_t *words = reinterpret_cast<const uint32_t *>(bytes);
    if (reinterpret_cast<uintptr_t>(bytes) % 4 != 0) {
        aligned_copy.resize(n_words);
        std::memcpy(aligned_copy.data(), bytes, n_words * 4);
        words = aligned_copy.data();
    }

    if (n_vec > 0) {
        const uint32_t grid = static_cast<uint32_t>(
            std::min<uint64_t>((n_vec + kSimpleHashThreads - 1) / kSimpleHashThreads, kSimpleHashMaxBlocks));
        // the device kernel does this arithmetic in 32 bits
        const uint32_t total_vec = static_cast<uint32_t>(n_vec);
        const uint32_t per_block = (total_vec + grid - 1) / grid;
        std::vector<uint32_t> partial(grid, 0);
        auto run_blocks = [&](uint32_t first, uint32_t last) {
            for (uint32_t b = first; b < last; ++b) {
                const uint64_t begin = static_cast<uint64_t>(b) * per_block;
                const uint64_t end = std::min<uint64_t>(begin + per_block, total_vec);
                partial[b] = begin < end ? block_partial(words, begin, end) : simplehash_block_fold(std::vector<uint32_t>(kSimpleHashThreads, 0).data());
            }
        };
        unsigned n_threads = 1;
        if (n_bytes >= (8u << 20)) n_threads = std::max(1u, std::min(16u, std::thread::hardware_concurrency() / 2));
        if (hash_thread_cap()) n_threads = std::min(n_threads, hash_thread_cap());
        if (n_threads <= 1 || grid < n_threads) {
            run_blocks(0, grid);
        } else {
            std::vector<std::thread> pool;
            for (unsigned t = 0; t < n_threads; ++t) {
                const uint32_t first = static_cast<uint32_t>(static_cast<uint64_t>(grid) * t / n_threads);
                const uint32_t last = static_cast<uint32_t>(static_cast<uint64_t>(grid) * (t + 1) / n_threads);
                pool.emplace_back(run_blocks, first, last);
            }
            for (auto &th : pool) th.join();
        }
        // final pass: 256 threads stride over the partials
        uint32_t per_thread[kSimpleHashThreads];
        for (uint32_t t = 0; t < kSimpleHashThreads; ++t) {
            uint32_t h = 0;
            for (uint32_t i = t; i < grid; i += kSimpleHashThreads) h = hash_combine(h, partial[i]);
            per_thread[t] = h;
        }
        result = simplehash_block_fold(per_thread);
    }
    for (size_t w = n_vec * 4; w < n_words; ++w) result = hash_combine(result, words[w]);
    if (const size_t tail = n_bytes % 4) {
        uint32_t last = 0;
        for (size_t i = 0; i < tail; ++i) last |= static_cast<uint32_t>(bytes[n_words * 4 + i]) << (8 * i);
        result = hash_combine(result, last);
    }
    return result;
}

} // namespace pccl::cpu
