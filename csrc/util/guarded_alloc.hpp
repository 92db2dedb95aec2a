// Guard-page allocations for hunting out-of-bounds accesses without a sanitizer build.
//
// guarded_alloc(n) returns n bytes whose last byte sits directly in front of an inaccessible page, and
// whose first byte is preceded by one (after at most 15 bytes of alignment slack, which is filled
// with a canary that guarded_free verifies).  Any overrun of a reduce, quantise or receive loop then
// faults at the offending instruction instead of corrupting a neighbour.
// The C++ end-to-end tests place the user buffers of the ring all-reduce in such blocks.
//
// Reference counterpart: ccoip/src/cpp/guarded_alloc.cpp:21-94 (compiled out there).
#pragma once
#include <cstddef>

namespace pccl {

void *guarded_alloc(size_t n_bytes);
// returns false when the canary in front of the block was overwritten
bool guarded_free(void *p);
size_t guarded_live_blocks();

} // namespace pccl
