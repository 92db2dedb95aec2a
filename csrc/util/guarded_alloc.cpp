#include "util/guarded_alloc.hpp"

#include <cstdint>
#include <cstring>
#include <mutex>
#include <unordered_map>

#include <sys/mman.h>
#include <unistd.h>

#include "util/log.hpp"

namespace pccl {
namespace {

struct Block {
    uint8_t *map;     // start of the mapping (first guard page)
    size_t map_bytes; // whole mapping
    size_t slack;     // canary bytes between the end of the leading guard and the user pointer
};
constexpr uint8_t kCanary = 0xA5;

std::mutex g_mu;
std::unordered_map<void *, Block> &blocks() {
    static auto *m = new std::unordered_map<void *, Block>(); // leaked on purpose: frees may run during exit
    return *m;
}

} // namespace

void *guarded_alloc(size_t n_bytes) {
    const size_t page = static_cast<size_t>(sysconf(_SC_PAGESIZE));
    const size_t want = n_bytes ? n_bytes : 1;
    const size_t body = (want + page - 1) / page * page;
    const size_t total = body + 2 * page;
    void *m = mmap(nullptr, total, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
    if (m == MAP_FAILED) return nullptr;
    auto *base = static_cast<uint8_t *>(m);
    mprotect(base, page, PROT_NONE);
    mprotect(base + page + body, page, PROT_NONE);
    // right-align so that the byte after the block is the trailing guard; keep 16-byte alignment for
    // vector loads by rounding the start down, which leaves < 16 bytes in front of the guard unused
    // only when n is not a multiple of 16 -- in that case the tail stays accessible but carries a canary
    uint8_t *user = base + page + body - want;
    user = reinterpret_cast<uint8_t *>(reinterpret_cast<uintptr_t>(user) & ~uintptr_t{15});
    const size_t slack = static_cast<size_t>(user - (base + page));
    std::memset(base + page, kCanary, slack);
    std::memset(user + want, kCanary, static_cast<size_t>(base + page + body - (user + want)));
    std::lock_guard<std::mutex> lock(g_mu);
    blocks()[user] = Block{base, total, slack};
    return user;
}

bool guarded_free(void *p) {
    if (!p) return true;
    Block b{};
    {
        std::lock_guard<std::mutex> lock(g_mu);
        auto it = blocks().find(p);
        if (it == blocks().end()) {
            PLOG(ERR) << "guarded_free of a pointer that guarded_alloc did not return";
            return false;
        }
        b = it->second;
        blocks().erase(it);
    }
    const size_t page = static_cast<size_t>(sysconf(_SC_PAGESIZE));
    bool intact = true;
    for (size_t i = 0; i < b.slack && intact; ++i) intact = b.map[page + i] == kCanary;
    if (!intact) PLOG(ERR) << "guarded block: bytes in front of the allocation were overwritten";
    munmap(b.map, b.map_bytes);
    return intact;
}

size_t guarded_live_blocks() {
    std::lock_guard<std::mutex> lock(g_mu);
    return blocks().size();
}

} // namespace pccl
