// Big-endian serialisation primitives of the CCoIP wire format (format spec: reference
// ccoip/internal/ccoip_packet_buffer.hpp:35-63,189-244).  Integers are written MSB first, floating
// point values are bit-cast to the same-width integer first, bool is one byte, strings are a u64
// length followed by raw bytes.
#pragma once
#include <bit>
#include <cstdint>
#include <cstring>
#include <span>
#include <string>
#include <type_traits>
#include <vector>

namespace pccl::wire {

class Writer {
public:
    Writer() = default;
    explicit Writer(size_t reserve) { buf_.reserve(reserve); }

    template <typename T> void put(T v) {
        if constexpr (std::is_same_v<T, bool>) {
            buf_.push_back(v ? 1 : 0);
        } else if constexpr (std::is_same_v<T, float>) {
            put<uint32_t>(std::bit_cast<uint32_t>(v));
        } else if constexpr (std::is_same_v<T, double>) {
            put<uint64_t>(std::bit_cast<uint64_t>(v));
        } else {
            static_assert(std::is_integral_v<T>);
            using U = std::make_unsigned_t<T>;
            const U u = static_cast<U>(v);
            for (int shift = (sizeof(U) - 1) * 8; shift >= 0; shift -= 8) buf_.push_back(static_cast<uint8_t>(u >> shift));
        }
    }
    void put_bytes(const void *p, size_t n) {
        const auto *b = static_cast<const uint8_t *>(p);
        buf_.insert(buf_.end(), b, b + n);
    }
    void put_string(const std::string &s) {
        put<uint64_t>(s.size());
        put_bytes(s.data(), s.size());
    }
    const std::vector<uint8_t> &bytes() const { return buf_; }
    std::vector<uint8_t> &bytes() { return buf_; }
    size_t size() const { return buf_.size(); }

private:
    std::vector<uint8_t> buf_;
};

// Reads never throw; a short buffer sets a sticky error flag and yields zeros.
class Reader {
public:
    Reader(const uint8_t *p, size_t n) : p_(p), n_(n) {}
    explicit Reader(std::span<const uint8_t> s) : p_(s.data()), n_(s.size()) {}

    template <typename T> T get() {
        if constexpr (std::is_same_v<T, bool>) {
            return get<uint8_t>() != 0;
        } else if constexpr (std::is_same_v<T, float>) {
            return std::bit_cast<float>(get<uint32_t>());
        } else if constexpr (std::is_same_v<T, double>) {
            return std::bit_cast<double>(get<uint64_t>());
        } else {
            using U = std::make_unsigned_t<T>;
            if (n_ - pos_ < sizeof(U)) {
                bad_ = true;
                pos_ = n_;
                return T{};
            }
            U u = 0;
            for (size_t i = 0; i < sizeof(U); ++i) u = static_cast<U>((u << 8) | p_[pos_ + i]);
            pos_ += sizeof(U);
            return static_cast<T>(u);
        }
    }
    bool get_bytes(void *dst, size_t n) {
        if (n_ - pos_ < n) {
            bad_ = true;
            pos_ = n_;
            return false;
        }
        std::memcpy(dst, p_ + pos_, n);
        pos_ += n;
        return true;
    }
    std::string get_string() {
        const uint64_t len = get<uint64_t>();
        if (bad_ || n_ - pos_ < len) {
            bad_ = true;
            pos_ = n_;
            return {};
        }
        std::string s(reinterpret_cast<const char *>(p_ + pos_), len);
        pos_ += len;
        return s;
    }
    // Guards element counts read from the wire: each element needs at least min_elem_bytes.
    bool plausible_count(uint64_t count, size_t min_elem_bytes) {
        if (bad_ || count > (n_ - pos_) / (min_elem_bytes ? min_elem_bytes : 1)) {
            bad_ = true;
            return false;
        }
        return true;
    }
    bool ok() const { return !bad_; }
    size_t remaining() const { return n_ - pos_; }
    const uint8_t *cursor() const { return p_ + pos_; }

private:
    const uint8_t *p_;
    size_t n_;
    size_t pos_ = 0;
    bool bad_ = false;
};

} // namespace pccl::wire
