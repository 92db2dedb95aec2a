// Environment-variable knobs (reference list: SURVEY §5.6) plus the device-plane knobs of this build.
#pragma once
#include <cstdint>
#include <cstdlib>
#include <string>

namespace pccl {

inline uint64_t env_u64(const char *name, uint64_t dflt) {
    const char *v = std::getenv(name);
    if (!v || !*v) return dflt;
    char *end = nullptr;
    const unsigned long long x = strtoull(v, &end, 10);
    if (end == v) return dflt;
    return x;
}
inline bool env_flag(const char *name, bool dflt) {
    const char *v = std::getenv(name);
    if (!v || !*v) return dflt;
    return !(v[0] == '0' || v[0] == 'n' || v[0] == 'N' || v[0] == 'f' || v[0] == 'F');
}

} // namespace pccl
