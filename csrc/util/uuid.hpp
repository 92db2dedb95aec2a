// 16-byte peer identity assigned by the master (reference: ccoip/internal/ccoip_types.hpp:100-144,
// ccoip/src/cpp/uuid_utils.cpp:6-20).
#pragma once
#include <array>
#include <cstdint>
#include <cstring>
#include <functional>
#include <random>
#include <string>

namespace pccl {

struct Uuid {
    std::array<uint8_t, 16> data{};
    bool operator==(const Uuid &o) const { return data == o.data; }
    bool operator!=(const Uuid &o) const { return data != o.data; }
    bool operator<(const Uuid &o) const { return data < o.data; } // byte-wise lexicographic
    std::string str() const {
        static const char *hex = "0123456789abcdef";
        std::string s;
        for (int i = 0; i < 16; ++i) {
            if (i == 4 || i == 6 || i == 8 || i == 10) s.push_back('-');
            s.push_back(hex[data[i] >> 4]);
            s.push_back(hex[data[i] & 15]);
        }
        return s;
    }
    std::string hex() const {
        static const char *digits = "0123456789abcdef";
        std::string s;
        for (int i = 0; i < 16; ++i) {
            s.push_back(digits[data[i] >> 4]);
            s.push_back(digits[data[i] & 15]);
        }
        return s;
    }
    static Uuid random_v4() {
        static thread_local std::mt19937_64 rng{std::random_device{}()};
        Uuid u;
        const uint64_t a = rng(), b = rng();
        std::memcpy(u.data.data(), &a, 8);
        std::memcpy(u.data.data() + 8, &b, 8);
        u.data[6] = static_cast<uint8_t>((u.data[6] & 0x0f) | 0x40);
        u.data[8] = static_cast<uint8_t>((u.data[8] & 0x3f) | 0x80);
        return u;
    }
};

} // namespace pccl

template <> struct std::hash<pccl::Uuid> {
    size_t operator()(const pccl::Uuid &u) const noexcept {
        uint64_t a, b;
        std::memcpy(&a, u.data.data(), 8);
        std::memcpy(&b, u.data.data() + 8, 8);
        return static_cast<size_t>(a * 0x9e3779b97f4a7c15ull ^ b);
    }
};
