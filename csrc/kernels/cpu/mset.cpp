// Host twin of the mset64 content hash (definition and rationale: csrc/kernels/mset_formula.hpp).  The terms are
// independent, so the loop is split over threads and the partial sums added -- any split gives the same value.
#include <algorithm>
#include <cstring>
#include <thread>
#include <vector>

#include "../mset_formula.hpp"
#include "hash.hpp"

namespace pccl::cpu {

uint64_t mset64_raw(const void *data, size_t n_bytes, uint64_t first_packet) {
    const auto *bytes = static_cast<const uint8_t *>(data);
    const size_t full = n_bytes / 16;
    auto span = [&](size_t lo, size_t hi) {
        uint64_t sum = 0;
        for (size_t i = lo; i < hi; ++i) {
            uint32_t w[4];
            std::memcpy(w, bytes + i * 16, 16);
            sum += mset::term(first_packet + i, w[0], w[1], w[2], w[3]);
        }
        return sum;
    };
    uint64_t raw = 0;
    const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
    size_t n_threads = std::min<size_t>(hw > 16 ? 16 : hw, full / (1u << 16)); // below 1 MiB per thread it is not worth it
    if (hash_thread_cap()) n_threads = std::min<size_t>(n_threads, hash_thread_cap());
    if (n_threads <= 1) {
        raw = span(0, full);
    } else {
        std::vector<uint64_t> part(n_threads, 0);
        std::vector<std::thread> pool;
        const size_t per = (full + n_threads - 1) / n_threads;
        for (size_t t = 0; t < n_threads; ++t)
            pool.emplace_back([&, t] { part[t] = span(std::min(full, t * per), std::min(full, (t + 1) * per)); });
        for (auto &th : pool) th.join();
        for (uint64_t p : part) raw += p;
    }
    if (const size_t tail = n_bytes % 16) {
        uint32_t w[4] = {0, 0, 0, 0};
        std::memcpy(w, bytes + full * 16, tail);
        raw += mset::term(first_packet + full, w[0], w[1], w[2], w[3]);
    }
    return raw;
}

uint64_t mset64(const void *data, size_t n_bytes) { return mset::finalize(mset64_raw(data, n_bytes, 0), n_bytes); }

} // namespace pccl::cpu
