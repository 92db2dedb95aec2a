// Device data plane of a communicator: everything that touches CUDA.
//
// The reference library has no device data plane (its all-reduce asserts host memory,
// python/framework/pccl/_pccl.py:617; CUDA is only used to hash and stage shared-state tensors,
// ccoip/src/cpp/ccoip_client_handler.cpp:537-550,1140-1155).  This class is what makes peers that share
// one NVLink/NVSwitch domain reduce on the GPUs directly:
//
//   * every communicator owns an "arena": device memory created with the CUDA VMM API and
//     exported as a POSIX file descriptor; peers on the same host fetch the descriptor over an
//     abstract unix-domain socket named after the peer UUID (SCM_RIGHTS) and map it;
//   * the arena starts with a flag page used for cross-GPU barriers (values are derived from the
//     master-issued sequence number, so they only ever grow and stale values are harmless),
//     followed by staging slots;
//   * all-reduce = one kernel launch per rank (csrc/kernels/cuda/allreduce_p2p.cu).
//
// Everything here degrades gracefully: without a GPU, or with peers on other hosts, the
// communicator uses the TCP ring and stages device buffers through pinned host memory.
#pragma once
#include <cstddef>
#include <cstdint>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../util/uuid.hpp"
#include "../wire/packets.hpp"
#include "pccl.h"

namespace pccl::device {

struct PointerInfo {
    bool is_device = false;
    int device = -1;
};
PointerInfo classify_pointer(const void *p);
bool cuda_runtime_usable(); // a driver and at least one device are present

// Fused DiLoCo outer step (fp32, AVG): the reduction's input is params - inner, its output is consumed by the
// Nesterov update of params/momentum, and inner receives the new parameters.
struct OuterStepArgs {
    float *params = nullptr;
    float *inner = nullptr;
    float *momentum = nullptr;
    float lr = 0.7f, mu = 0.9f;
    bool first = false; // the momentum buffer holds nothing yet
};

struct DeviceReduceArgs {
    const void *src = nullptr;
    void *dst = nullptr;
    uint64_t count = 0;
    wire::DType dtype = wire::DType::F32;
    wire::RedOp op = wire::RedOp::Sum;
    wire::QuantAlgo quant = wire::QuantAlgo::None;
    wire::DType quant_dtype = wire::DType::U8;
    uint64_t seq = 0;
    uint32_t rank = 0, world = 1;
    bool use_nvls = false; // all peers agreed (through the master) to run this one over NVSwitch multicast
    const OuterStepArgs *outer = nullptr; // non-null: src/dst are ignored, see OuterStepArgs
    // lease path (no vote at the master): the operation is identified by (tag, lease_k) in the same-box sequencer
    bool leased = false;
    uint64_t tag = 0, lease_k = 0;
    // voted operations: every peer of the group holds the same-box sequencer (agreed through the master, like the plane), so
    // the operation takes its place and its resources from the common log; otherwise it runs in Commence order
    bool sequenced = false;
    bool want_hash = false; // produce the mset64 digest(s) of the output as a by-product
    void *user_stream = nullptr; // non-null: order the operation after this CUDA stream instead of the legacy default stream
    bool enqueue_only = false;   // return after the launch; the caller finishes with Plane::finish()
};

struct DeviceReduceResult {
    bool ok = false;
    bool aborted = false;
    pcclExtDataPlane_t plane = pcclExtPlaneNone;
    float device_ms = 0;
    uint32_t kernel_launches = 0;
    uint64_t tx_bytes = 0, rx_bytes = 0;
    bool revoked = false;         // gave up because the lease ended underneath it (not a data-plane failure)
    bool input_destroyed = false; // in place on symmetric memory: a failed attempt may have overwritten the input
    bool has_hash = false;
    uint64_t hash[2] = {0, 0};    // mset64 of the output (outer step: [0] parameters, [1] momentum)
    uint64_t phase_ns[8] = {0};   // GPU global timer at the kernel's phase boundaries (block 0)
    bool pending = false;         // enqueue_only: launched, not finished
    uint64_t ticket = 0;          // identifies the pending operation for finish()
};

class Plane {
public:
    static std::unique_ptr<Plane> create(); // never null; may be an inert object
    virtual ~Plane() = default;

    // Binds the plane to a peer identity and opens the rendezvous socket.
    virtual void start(const Uuid &self) = 0;
    virtual void stop() = 0;

    // Called after every successful topology round with the new ring of the caller's group.
    // Maps the arenas of all ring members that live on this host.
    virtual void on_ring_changed(const std::vector<Uuid> &ring) = 0;

    // What this peer could do for the described reduction: kCapP2P when the buffers are on the
    // GPU, every ring member is mapped and the operator/type is supported by the peer-to-peer
    // kernels; additionally kCapNvls when the multicast view exists and multimem can reduce it.
    static constexpr uint32_t kCapP2P = 1, kCapNvls = 2;
    virtual uint32_t capabilities(const DeviceReduceArgs &args) = 0;

    // Lease support: true when the same-box sequencer of the current ring is attached (operations can be ordered
    // without the master); multicast_ready() tells whether the NVSwitch view of the current ring exists here.
    virtual bool lease_ready() = 0;
    virtual bool multicast_ready() = 0;
    // Whether the multicast kernels can run this reduction (type, operator, size, ring size) -- the same answer on
    // every peer; combined with "every peer has the multicast view" it decides the plane of a leased operation.
    virtual bool nvls_eligible(const DeviceReduceArgs &args) = 0;
    // The lease ended (tag, or everything when all is set): leased operations in flight give up.
    virtual void revoke_leased(uint64_t tag, bool all) = 0;
    // Second half of an enqueue_only reduction: waits for the kernel and reports its verdict.
    virtual DeviceReduceResult finish(uint64_t ticket, const std::function<bool()> &abort_requested) = 0;

    // Device operations must run in the same order on every peer, otherwise two peers can each
    // run a different operation first (holding the staging slot and the SMs) and wait for each
    // other forever.
    // The order everybody agrees on is the order in which the master's Commence packets arrive:
    // `announce` is called for every device-plane Commence, in arrival order; `reduce` then waits
    // for its turn; an announced operation that is not going to run must be `cancel`led.
    virtual void announce(uint64_t seq) = 0;
    virtual void cancel(uint64_t seq) = 0;

    // Runs the reduction; `abort_requested` is polled while the kernel runs.
    virtual DeviceReduceResult reduce(const DeviceReduceArgs &args, const std::function<bool()> &abort_requested) = 0;

    // Stream-ordered staging helpers for the TCP fallback and for shared state.
    virtual bool copy_to_host(void *host_dst, const void *dev_src, size_t bytes) = 0;
    virtual bool copy_to_device(void *dev_dst, const void *host_src, size_t bytes) = 0;
    // Content hash of a device buffer; hash_type as on the wire.
    virtual bool hash_device(const void *dev_ptr, size_t bytes, wire::HashType type, uint64_t &out) = 0;
    // Many buffers at once: the kernels are queued back to back and waited for once (a model's parameter list is hundreds
    // of tensors; one synchronisation per tensor costs more than the hashing).
    struct HashJob {
        const void *ptr;
        size_t bytes;
        uint64_t result;
    };
    virtual bool hash_device_many(std::vector<HashJob> &jobs, wire::HashType type) = 0;

    // Shared-state transfer between peers of one NVLink domain: the distributor stages a chunk in its
    // arena (export), the requester copies it out of the mapped arena (import).  `exchange_bytes` is
    // the chunk capacity, 0 when the plane has no arena; `exchange_lock` serialises users of the
    // staging area on the distributor.
    virtual bool peer_mapped(const Uuid &peer) = 0;
    virtual size_t exchange_bytes() = 0;
    virtual std::unique_lock<std::mutex> exchange_lock() = 0;
    // `slot` < exchange_slots(): with two of them the distributor stages chunk i+1 while the requester still reads chunk i.
    virtual size_t exchange_slots() = 0;
    virtual bool export_chunk(size_t slot, const void *src, bool src_on_device, size_t bytes) = 0;
    virtual bool import_chunk(size_t slot, const Uuid &from, void *dst, bool dst_on_device, size_t bytes) = 0;
    // The same in two halves, so that the copies of consecutive chunks queue up back to back: begin() enqueues the copy,
    // finish_oldest() waits for the oldest one still pending (false when none is, or on failure).
    virtual bool import_chunk_begin(size_t slot, const Uuid &from, void *dst, bool dst_on_device, size_t bytes) = 0;
    virtual bool import_chunk_finish_oldest() = 0;

    // Fusing many small reductions into one: a reusable device buffer (from the symmetric heap when
    // it fits, so the fused reduction needs no staging) and batched device-to-device copies.
    struct Part {
        void *ptr;
        size_t bytes;
    };
    // The buffer is one per plane: whoever uses it holds `fuse_lock()` from fuse_buffer() until it is done with the contents.
    // The holders are synchronous collective calls (fused tensor lists, two-pass outer step) that keep the lock across
    // their master vote; like all collectives they have to be issued in the same order on every peer -- two threads
    // that start them in opposite orders on two peers would wait for each other's vote forever.
    virtual void *fuse_buffer(size_t bytes) = 0;
    virtual std::unique_lock<std::mutex> fuse_lock() = 0;
    virtual bool gather(const std::vector<Part> &parts, void *fused, bool into_fused) = 0;

    // The two halves of the outer step as separate passes over device memory (used when the fused kernel cannot
    // run: TCP plane, operands larger than a staging slot).  Synchronous.
    virtual bool pseudo_grad(float *out, const float *params, const float *inner, size_t count) = 0;
    virtual bool outer_nesterov(const OuterStepArgs &step, const float *grad, size_t count) = 0;

    // Page-locked host memory for staging (plain malloc when there is no GPU).
    virtual void *host_alloc(size_t bytes) = 0;
    virtual void host_free(void *ptr) = 0;

    virtual void *symmetric_alloc(size_t bytes) = 0;
    virtual bool symmetric_free(void *ptr) = 0;
};

} // namespace pccl::device
