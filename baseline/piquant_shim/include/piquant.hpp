// Minimal stand-in for PrimeIntellect's piquant library, which the reference includes
// unconditionally (ccoip/internal/piquant_utils.hpp:5) but which is missing from the scraped tree.
// It exists only so that the UNMODIFIED reference sources configure and compile for the baseline
// arm of bench.py.  The timed baseline configurations use pcclQuantNone / pcclQuantMinMax, which
// never reach this code.  Plain zero-point/scale affine quantisation, single threaded.
#pragma once
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <limits>
#include <span>
#include <stdexcept>
#include <utility>

namespace piquant {

enum class dtype { uint8, int8, uint16, int16, uint32, int32, uint64, int64, f32, f64 };
enum class round_mode { nearest, stochastic };
enum class reduce_op { set, add };

namespace detail {
template <typename F> void with_int(dtype t, F &&f) {
    switch (t) {
    case dtype::uint8: f(std::uint8_t{}); break;
    case dtype::int8: f(std::int8_t{}); break;
    case dtype::uint16: f(std::uint16_t{}); break;
    case dtype::int16: f(std::int16_t{}); break;
    case dtype::uint32: f(std::uint32_t{}); break;
    case dtype::int32: f(std::int32_t{}); break;
    case dtype::uint64: f(std::uint64_t{}); break;
    case dtype::int64: f(std::int64_t{}); break;
    default: throw std::logic_error{"piquant shim: integer type expected"};
    }
}
template <typename F> void with_float(dtype t, F &&f) {
    switch (t) {
    case dtype::f32: f(float{}); break;
    case dtype::f64: f(double{}); break;
    default: throw std::logic_error{"piquant shim: float type expected"};
    }
}
} // namespace detail

class context {
public:
    explicit context(std::size_t) {}

    template <typename T> std::pair<float, std::int64_t> compute_quant_config_from_data(std::span<const T> data, dtype q) {
        double lo = 0, hi = 0;
        if (!data.empty()) {
            lo = hi = static_cast<double>(data[0]);
            for (const T v : data) {
                lo = std::min(lo, static_cast<double>(v));
                hi = std::max(hi, static_cast<double>(v));
            }
        }
        double qlo = 0, qhi = 255;
        detail::with_int(q, [&](auto tag) {
            using Q = decltype(tag);
            qlo = static_cast<double>(std::numeric_limits<Q>::lowest());
            qhi = static_cast<double>(std::numeric_limits<Q>::max());
        });
        if (hi == lo) return {lo != 0 ? static_cast<float>(lo) : 1.0f, 0};
        const float scale = static_cast<float>((hi - lo) / (qhi - qlo));
        return {scale, static_cast<std::int64_t>(std::nearbyint(qlo - lo / scale))};
    }

    template <typename In, typename Out>
    void quantize(In in, dtype in_t, Out out, dtype out_t, float scale, std::int64_t zp, round_mode) {
        detail::with_float(in_t, [&](auto ftag) {
            using F = decltype(ftag);
            detail::with_int(out_t, [&](auto qtag) {
                using Q = decltype(qtag);
                const auto *x = reinterpret_cast<const F *>(in.data());
                auto *q = reinterpret_cast<Q *>(out.data());
                const std::size_t n = in.size_bytes() / sizeof(F);
                for (std::size_t i = 0; i < n; ++i) {
                    double v = std::nearbyint(static_cast<double>(x[i]) / scale) + static_cast<double>(zp);
                    v = std::min(std::max(v, static_cast<double>(std::numeric_limits<Q>::lowest())), static_cast<double>(std::numeric_limits<Q>::max()));
                    q[i] = static_cast<Q>(v);
                }
            });
        });
    }

    template <typename In, typename Out>
    void dequantize(In in, dtype in_t, Out out, dtype out_t, float scale, std::int64_t zp, reduce_op op) {
        detail::with_int(in_t, [&](auto qtag) {
            using Q = decltype(qtag);
            detail::with_float(out_t, [&](auto ftag) {
                using F = decltype(ftag);
                const auto *q = reinterpret_cast<const Q *>(in.data());
                auto *x = reinterpret_cast<F *>(out.data());
                const std::size_t n = in.size_bytes() / sizeof(Q);
                for (std::size_t i = 0; i < n; ++i) {
                    const F v = static_cast<F>((static_cast<double>(q[i]) - static_cast<double>(zp)) * scale);
                    x[i] = op == reduce_op::add ? static_cast<F>(x[i] + v) : v;
                }
            });
        });
    }

    // in -> quantise -> dequantise -> out (both float, `mid` is the code type)
    template <typename In, typename Out>
    void quantize_dequantize_fused(In in, dtype in_t, Out out, dtype mid, float scale, std::int64_t zp, round_mode, reduce_op op) {
        detail::with_float(in_t, [&](auto ftag) {
            using F = decltype(ftag);
            detail::with_int(mid, [&](auto qtag) {
                using Q = decltype(qtag);
                const auto *x = reinterpret_cast<const F *>(in.data());
                auto *y = reinterpret_cast<F *>(out.data());
                const std::size_t n = in.size_bytes() / sizeof(F);
                for (std::size_t i = 0; i < n; ++i) {
                    double v = std::nearbyint(static_cast<double>(x[i]) / scale) + static_cast<double>(zp);
                    v = std::min(std::max(v, static_cast<double>(std::numeric_limits<Q>::lowest())), static_cast<double>(std::numeric_limits<Q>::max()));
                    const F d = static_cast<F>((static_cast<double>(static_cast<Q>(v)) - static_cast<double>(zp)) * scale);
                    y[i] = op == reduce_op::add ? static_cast<F>(y[i] + d) : d;
                }
            });
        });
    }
};

} // namespace piquant
