// Prologue and epilogue of the fused DiLoCo outer step (see allreduce_p2p.cu / allreduce_nvls.cu):
//   prologue  pseudo-gradient packet  d = params - inner           (instead of "copy the input into the arena")
//   epilogue  Nesterov-SGD on the averaged pseudo-gradient g       (instead of "copy the arena into the output")
//             buf = first ? g : mu*buf + g;  params -= lr*(g + mu*buf);  inner = params
// The arithmetic is the one of pseudo_grad_kernel / outer_nesterov_kernel in elementwise.cu (torch.optim.SGD with
// nesterov=True), which in turn restates what the reference's example does with torch ops on the host
// (python/examples/nanogpt_diloco/sync_diloco.py:480-530).
#pragma once
#include "common.cuh"
#include "hash_mset.cuh"
#include "pccl_kernels.h"

namespace pccl::cuda {

struct OuterStepDev {
    const float *inner_in; // parameters after the inner steps
    float *params;         // outer parameters, updated by the epilogue
    float *momentum;       // Nesterov buffer, updated by the epilogue
    float *inner_out;      // receives the new parameters (usually == inner_in); may be null
    float lr, mu;
    int first; // momentum buffer holds nothing yet
    int aligned; // all four pointers are 16-byte aligned
};

inline OuterStepDev make_outer_step_dev(const pcclKernelOuterStep_t &step) {
    OuterStepDev s{};
    s.inner_in = step.inner_in;
    s.params = step.params;
    s.momentum = step.momentum;
    s.inner_out = step.inner_out;
    s.lr = step.lr;
    s.mu = step.mu;
    s.first = step.first;
    s.aligned = ((reinterpret_cast<uintptr_t>(s.inner_in) | reinterpret_cast<uintptr_t>(s.params) | reinterpret_cast<uintptr_t>(s.momentum) |
                  reinterpret_cast<uintptr_t>(s.inner_out)) & 15) == 0;
    return s;
}

__device__ __forceinline__ uint4 outer_delta_packet(const OuterStepDev &s, uint64_t v, uint64_t count) {
    const uint64_t e = v * 4;
    if (s.aligned && e + 4 <= count) {
        const uint4 a = ld_stream_16(s.params + e), b = ld_stream_16(s.inner_in + e);
        uint4 r;
        r.x = __float_as_uint(__uint_as_float(a.x) - __uint_as_float(b.x));
        r.y = __float_as_uint(__uint_as_float(a.y) - __uint_as_float(b.y));
        r.z = __float_as_uint(__uint_as_float(a.z) - __uint_as_float(b.z));
        r.w = __float_as_uint(__uint_as_float(a.w) - __uint_as_float(b.w));
        return r;
    }
    uint32_t r[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) r[i] = e + i < count ? __float_as_uint(s.params[e + i] - s.inner_in[e + i]) : 0u;
    return make_uint4(r[0], r[1], r[2], r[3]);
}

__device__ __forceinline__ void outer_update_one(const OuterStepDev &s, float g, float p_old, float m_old, float &p_new, float &m_new) {
    m_new = s.first ? g : fmaf(s.mu, m_old, g);
    p_new = p_old - s.lr * fmaf(s.mu, m_new, g);
}

// loads of the epilogue, split from the arithmetic so that callers can keep several packets in flight
struct OuterLoaded {
    uint4 p, m;
};
__device__ __forceinline__ OuterLoaded outer_load_packet(const OuterStepDev &s, uint64_t v, uint64_t count) {
    OuterLoaded l{};
    const uint64_t e = v * 4;
    if (s.aligned && e + 4 <= count) {
        l.p = ld_stream_16(s.params + e);
        if (!s.first) l.m = ld_stream_16(s.momentum + e);
    }
    return l;
}

// dig_p / dig_m (optional): mset64 accumulators of the new parameters and of the new momentum buffer, so that the content
// hashes the next shared-state synchronisation needs are a by-product of this pass
__device__ __forceinline__ void outer_apply_packet(const OuterStepDev &s, uint64_t v, uint64_t count, uint4 g, const OuterLoaded &l,
                                                   MsetAcc *dig_p = nullptr, MsetAcc *dig_m = nullptr) {
    const uint64_t e = v * 4;
    if (s.aligned && e + 4 <= count) {
        float p[4], m[4];
        outer_update_one(s, __uint_as_float(g.x), __uint_as_float(l.p.x), __uint_as_float(l.m.x), p[0], m[0]);
        outer_update_one(s, __uint_as_float(g.y), __uint_as_float(l.p.y), __uint_as_float(l.m.y), p[1], m[1]);
        outer_update_one(s, __uint_as_float(g.z), __uint_as_float(l.p.z), __uint_as_float(l.m.z), p[2], m[2]);
        outer_update_one(s, __uint_as_float(g.w), __uint_as_float(l.p.w), __uint_as_float(l.m.w), p[3], m[3]);
        const uint4 pv = make_uint4(__float_as_uint(p[0]), __float_as_uint(p[1]), __float_as_uint(p[2]), __float_as_uint(p[3]));
        const uint4 mv = make_uint4(__float_as_uint(m[0]), __float_as_uint(m[1]), __float_as_uint(m[2]), __float_as_uint(m[3]));
        st_stream_16(s.momentum + e, mv);
        st_stream_16(s.params + e, pv);
        if (s.inner_out) st_stream_16(s.inner_out + e, pv);
        if (dig_p) dig_p->add(v, pv, count * 4);
        if (dig_m) dig_m->add(v, mv, count * 4);
        return;
    }
    const uint32_t gw[4] = {g.x, g.y, g.z, g.w};
    uint32_t pw[4] = {0, 0, 0, 0}, mw[4] = {0, 0, 0, 0};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        if (e + i >= count) break;
        float p, m;
        outer_update_one(s, __uint_as_float(gw[i]), s.params[e + i], s.first ? 0.0f : s.momentum[e + i], p, m);
        s.momentum[e + i] = m;
        s.params[e + i] = p;
        if (s.inner_out) s.inner_out[e + i] = p;
        pw[i] = __float_as_uint(p);
        mw[i] = __float_as_uint(m);
    }
    if (dig_p) dig_p->add(v, make_uint4(pw[0], pw[1], pw[2], pw[3]), count * 4);
    if (dig_m) dig_m->add(v, make_uint4(mw[0], mw[1], mw[2], mw[3]), count * 4);
}

} // namespace pccl::cuda
