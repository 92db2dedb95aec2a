// Host-side element-wise kernels of the TCP data plane and the CPU oracle for the GPU kernels.
// Semantics follow the reference (ccoip/src/cpp/reduce_kernels.cpp:56-215 for the operators,
// :437-510 for AVG finalisation): AVG accumulates like SUM and is divided by the world size at the
// end, in the element type (integer division for integer types).
#pragma once
#include <cstddef>
#include <cstdint>

#include "../../wire/packets.hpp"

namespace pccl::cpu {

// dst[i] = dst[i] (op) src[i]; Set copies, Avg behaves like Sum.
void reduce_inplace(void *dst, const void *src, size_t count, wire::DType dtype, wire::RedOp op);
// dst[i] /= world (only meaningful for Avg)
void finalize_avg(void *dst, size_t count, wire::DType dtype, uint64_t world);
// Left fold helper used by tests: out = in[0] op in[1] op ... (each `count` elements)
bool dtype_supported(wire::DType dtype);

} // namespace pccl::cpu
