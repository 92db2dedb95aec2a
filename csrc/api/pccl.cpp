// C API: argument validation, enum translation, thin forwarding to client::Client / master::Master.
// Error-code mapping follows the reference (src/pccl.cpp:85-116 connect, :196-241 topology, :246-343
// reductions, :345-523 multiple-with-retry, :545-620 shared state, :622-681 master).
#include "pccl.h"

#include <atomic>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <unordered_set>
#include <vector>

#include "../client/client.hpp"
#include "../device/plane.hpp"
#include "../kernels/cpu/hash.hpp"
#include "../kernels/cpu/quantize.hpp"
#include "../master/master.hpp"
#include "../util/log.hpp"

using namespace pccl;

struct pcclComm_t {
    pcclCommCreateParams_t params{};
    std::unique_ptr<client::Client> client;
    // scratch of the fused form of pcclAllReduceMultipleWithRetry on host tensors: kept between calls (a training loop makes
    // the same call every step; a fresh half-gigabyte vector would be zero-filled and page-faulted in every time)
    std::mutex fuse_mu;
    std::unique_ptr<uint8_t[]> fuse_buf;
    size_t fuse_bytes = 0;
};

struct pcclMasterInstanceState_t {
    std::unique_ptr<master::Master> master;
    bool running = false, interrupted = false, awaited = false;
};

static std::atomic<bool> g_initialised{false};

#define PCCL_REQUIRE_INIT()                                                                                            \
    if (!g_initialised.load()) return pcclNotInitialized
#define PCCL_REQUIRE(cond, code)                                                                                       \
    if (!(cond)) return code

// pccl and wire number the data types differently (reference: src/pccl.cpp:123-152)
static bool to_wire_dtype(pcclDataType_t t, wire::DType &out) {
    switch (t) {
    case pcclUint8: out = wire::DType::U8; return true;
    case pcclInt8: out = wire::DType::I8; return true;
    case pcclInt16: out = wire::DType::I16; return true;
    case pcclUint16: out = wire::DType::U16; return true;
    case pcclUint32: out = wire::DType::U32; return true;
    case pcclInt32: out = wire::DType::I32; return true;
    case pcclUint64: out = wire::DType::U64; return true;
    case pcclInt64: out = wire::DType::I64; return true;
    case pcclFloat16: out = wire::DType::F16; return true;
    case pcclBFloat16: out = wire::DType::BF16; return true;
    case pcclFloat: out = wire::DType::F32; return true;
    case pcclDouble: out = wire::DType::F64; return true;
    case pcclFloat8E4M3: out = wire::DType::F8E4M3; return true;
    }
    return false;
}

static bool to_wire_op(pcclRedOp_t op, wire::RedOp &out) {
    switch (op) {
    case pcclSum: out = wire::RedOp::Sum; return true;
    case pcclAvg: out = wire::RedOp::Avg; return true;
    case pcclProd: out = wire::RedOp::Prod; return true;
    case pcclMax: out = wire::RedOp::Max; return true;
    case pcclMin: out = wire::RedOp::Min; return true;
    }
    return false;
}

static bool to_wire_quant(pcclQuantizationAlgorithm_t q, wire::QuantAlgo &out) {
    switch (q) {
    case pcclQuantNone: out = wire::QuantAlgo::None; return true;
    case pcclQuantMinMax: out = wire::QuantAlgo::MinMax; return true;
    case pcclQuantZeroPointScale: out = wire::QuantAlgo::ZeroPointScale; return true;
    case pcclQuantBlockMinMax: out = wire::QuantAlgo::BlockMinMax; return true;
    case pcclQuantBlockFp8: out = wire::QuantAlgo::BlockFp8; return true;
    }
    return false;
}

static pcclResult_t build_request(const void *send, void *recv, const pcclReduceDescriptor_t *d, client::ReduceRequest &r) {
    PCCL_REQUIRE(send && recv && d, pcclInvalidArgument);
    PCCL_REQUIRE(d->count > 0, pcclInvalidArgument);
    r.send = send;
    r.recv = recv;
    r.count = d->count;
    r.tag = d->tag;
    PCCL_REQUIRE(to_wire_dtype(d->src_descriptor.datatype, r.dtype) && r.dtype != wire::DType::F8E4M3, pcclInvalidArgument);
    PCCL_REQUIRE(to_wire_op(d->op, r.op), pcclInvalidArgument);
    PCCL_REQUIRE(to_wire_quant(d->quantization_options.algorithm, r.quant), pcclInvalidArgument);
    if (r.quant != wire::QuantAlgo::None) {
        PCCL_REQUIRE(to_wire_dtype(d->quantization_options.quantized_datatype, r.quant_dtype), pcclInvalidArgument);
        PCCL_REQUIRE(cpu::quant_supported(r.quant, r.dtype, r.quant_dtype), pcclInvalidArgument);
    }
    return pcclSuccess;
}

extern "C" {

pcclResult_t pcclInit(void) {
    g_initialised = true;
    return pcclSuccess;
}

size_t pcclExtDataTypeSize(pcclDataType_t datatype) {
    wire::DType t;
    return to_wire_dtype(datatype, t) ? wire::dtype_size(t) : 0;
}

pcclResult_t pcclGetBuildInfo(pcclBuildInfo_t *info) {
    PCCL_REQUIRE(info, pcclInvalidArgument);
#ifdef PCCL_NO_CUDA
    info->has_cuda_support = false;
#else
    info->has_cuda_support = true;
#endif
    return pcclSuccess;
}

pcclResult_t pcclCreateCommunicator(const pcclCommCreateParams_t *params, pcclComm_t **comm_out) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(params && comm_out, pcclInvalidArgument);
    auto *c = new pcclComm_t;
    c->params = *params;
    *comm_out = c;
    return pcclSuccess;
}

pcclResult_t pcclGetAttribute(const pcclComm_t *c, pcclAttribute_t attribute, int *out) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(c && out, pcclInvalidArgument);
    PCCL_REQUIRE(c->client, pcclInvalidUsage);
    switch (attribute) {
    case PCCL_ATTRIBUTE_GLOBAL_WORLD_SIZE: *out = static_cast<int>(c->client->global_world()); break;
    case PCCL_ATTRIBUTE_PEER_GROUP_WORLD_SIZE: *out = static_cast<int>(c->client->local_world()); break;
    case PCCL_ATTRIBUTE_NUM_DISTINCT_PEER_GROUPS: *out = static_cast<int>(c->client->num_groups()); break;
    case PCCL_ATTRIBUTE_LARGEST_PEER_GROUP_WORLD_SIZE: *out = static_cast<int>(c->client->largest_group()); break;
    default: return pcclInvalidArgument;
    }
    return pcclSuccess;
}

pcclResult_t pcclDestroyCommunicator(pcclComm_t *c) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(c, pcclInvalidArgument);
    if (c->client) {
        c->client->interrupt();
        c->client->join();
    }
    delete c;
    return pcclSuccess;
}

pcclResult_t pcclConnect(pcclComm_t *c) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(c, pcclInvalidArgument);
    PCCL_REQUIRE(!c->client, pcclInvalidUsage);
    client::ClientConfig cfg;
    cfg.master = c->params.master_address;
    cfg.peer_group = c->params.peer_group;
    cfg.pool_size = c->params.p2p_connection_pool_size ? c->params.p2p_connection_pool_size : 1;
    cfg.explicit_addresses = c->params.use_explicit_p2p_addresses;
    cfg.adv_p2p = c->params.advertised_p2p_address;
    cfg.adv_ss = c->params.advertised_shared_state_address;
    cfg.adv_bm = c->params.advertised_benchmark_address;
    if (c->params.internal_p2p_listen_port) cfg.p2p_port = c->params.internal_p2p_listen_port;
    if (c->params.internal_shared_state_listen_port) cfg.ss_port = c->params.internal_shared_state_listen_port;
    if (c->params.internal_benchmark_listen_port) cfg.bm_port = c->params.internal_benchmark_listen_port;
    c->client = std::make_unique<client::Client>(cfg);
    if (!c->client->connect()) {
        c->client->interrupt();
        c->client->join();
        c->client.reset();
        return pcclMasterConnectionFailed;
    }
    return pcclSuccess;
}

pcclResult_t pcclUpdateTopology(pcclComm_t *c) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(c, pcclInvalidArgument);
    PCCL_REQUIRE(c->client, pcclInvalidUsage);
    PCCL_REQUIRE(!c->client->any_reduce_running(), pcclPendingAsyncOps);
    return c->client->update_topology() ? pcclSuccess : pcclUpdateTopologyFailed;
}

pcclResult_t pcclArePeersPending(const pcclComm_t *c, bool *pending_out) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(c && pending_out, pcclInvalidArgument);
    PCCL_REQUIRE(c->client, pcclInvalidUsage);
    bool pending = false;
    if (!c->client->are_peers_pending(pending)) return pcclInternalError;
    *pending_out = pending;
    return pcclSuccess;
}

pcclResult_t pcclOptimizeTopology(const pcclComm_t *c) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(c, pcclInvalidArgument);
    PCCL_REQUIRE(c->client, pcclInvalidUsage);
    PCCL_REQUIRE(c->client->global_world() > 1, pcclInvalidUsage);
    PCCL_REQUIRE(!c->client->any_reduce_running(), pcclPendingAsyncOps);
    return c->client->optimize_topology() ? pcclSuccess : pcclTopologyOptimizationFailed;
}

pcclResult_t pcclAllReduceAsync(const void *sendbuff, void *recvbuff, const pcclReduceDescriptor_t *descriptor,
                                const pcclComm_t *c, pcclAsyncReduceOp_t *handle_out) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(c && handle_out, pcclInvalidArgument);
    PCCL_REQUIRE(c->client, pcclInvalidUsage);
    client::ReduceRequest req;
    if (const pcclResult_t rc = build_request(sendbuff, recvbuff, descriptor, req); rc != pcclSuccess) return rc;
    PCCL_REQUIRE(c->client->local_world() >= 2, pcclTooFewPeers);
    switch (c->client->all_reduce_async(req)) {
    case client::Client::Launch::Ok: break;
    case client::Client::Launch::DuplicateTag: return pcclInvalidArgument;
    case client::Client::Launch::NotConnected: return pcclInvalidUsage;
    }
    handle_out->comm = const_cast<pcclComm_t *>(c);
    handle_out->tag = descriptor->tag;
    return pcclSuccess;
}

pcclResult_t pcclAwaitAsyncReduce(const pcclAsyncReduceOp_t *handle, pcclReduceInfo_t *info_out) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(handle && handle->comm && handle->comm->client, pcclInvalidArgument);
    client::ReduceOutcome out;
    const bool ok = handle->comm->client->await_reduce(handle->tag, out);
    if (info_out) {
        info_out->local_world_size = out.world;
        info_out->tx_bytes = out.tx_bytes;
        info_out->rx_bytes = out.rx_bytes;
    }
    return ok ? pcclSuccess : pcclRankConnectionLost;
}

pcclResult_t pcclAllReduce(const void *sendbuff, void *recvbuff, const pcclReduceDescriptor_t *descriptor,
                           const pcclComm_t *c, pcclReduceInfo_t *info_out) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(c, pcclInvalidArgument);
    PCCL_REQUIRE(c->client, pcclInvalidUsage);
    client::ReduceRequest req;
    if (const pcclResult_t rc = build_request(sendbuff, recvbuff, descriptor, req); rc != pcclSuccess) return rc;
    PCCL_REQUIRE(c->client->local_world() >= 2, pcclTooFewPeers);
    client::ReduceOutcome out;
    switch (c->client->all_reduce_sync(req, out)) {
    case client::Client::Launch::Ok: break;
    case client::Client::Launch::DuplicateTag: return pcclInvalidArgument;
    case client::Client::Launch::NotConnected: return pcclInvalidUsage;
    }
    if (info_out) {
        info_out->local_world_size = out.world;
        info_out->tx_bytes = out.tx_bytes;
        info_out->rx_bytes = out.rx_bytes;
    }
    return out.ok ? pcclSuccess : pcclRankConnectionLost;
}

pcclResult_t pcclExtAllReduceOnStream(const void *sendbuff, void *recvbuff, const pcclReduceDescriptor_t *descriptor, const pcclComm_t *c,
                                      void *cuda_stream, int enqueue, pcclReduceInfo_t *info_out) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(c, pcclInvalidArgument);
    PCCL_REQUIRE(c->client, pcclInvalidUsage);
    client::ReduceRequest req;
    if (const pcclResult_t rc = build_request(sendbuff, recvbuff, descriptor, req); rc != pcclSuccess) return rc;
    PCCL_REQUIRE(device::classify_pointer(sendbuff).is_device && device::classify_pointer(recvbuff).is_device, pcclInvalidArgument);
    PCCL_REQUIRE(c->client->local_world() >= 2, pcclTooFewPeers);
    req.user_stream = cuda_stream ? cuda_stream : reinterpret_cast<void *>(uintptr_t(1)); // null = the legacy default stream (handle 1)
    req.enqueue = enqueue != 0;
    client::ReduceOutcome out;
    switch (c->client->all_reduce_on_stream(req, out)) {
    case client::Client::Launch::Ok: break;
    case client::Client::Launch::DuplicateTag: return pcclInvalidArgument;
    case client::Client::Launch::NotConnected: return pcclInvalidUsage;
    }
    if (info_out) *info_out = {out.world, out.tx_bytes, out.rx_bytes};
    return out.ok ? pcclSuccess : pcclRankConnectionLost;
}

pcclResult_t pcclExtTrustContentHashes(pcclComm_t *c, int on) {
    PCCL_REQUIRE(c, pcclInvalidArgument);
    PCCL_REQUIRE(c->client, pcclInvalidUsage);
    c->client->trust_content_hashes(on != 0);
    return pcclSuccess;
}

pcclResult_t pcclExtHashKernelsSkipped(const pcclComm_t *c, uint64_t *count_out) {
    PCCL_REQUIRE(c && count_out, pcclInvalidArgument);
    PCCL_REQUIRE(c->client, pcclInvalidUsage);
    *count_out = c->client->hash_kernels_skipped();
    return pcclSuccess;
}

pcclResult_t pcclExtWaitStreamOps(const pcclComm_t *c, uint64_t tag, int all_tags, pcclReduceInfo_t *info_out) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(c, pcclInvalidArgument);
    PCCL_REQUIRE(c->client, pcclInvalidUsage);
    client::ReduceOutcome out;
    const bool ok = c->client->wait_stream_ops(tag, all_tags != 0, out);
    if (info_out) *info_out = {out.world, out.tx_bytes, out.rx_bytes};
    return ok ? pcclSuccess : pcclRankConnectionLost;
}

pcclResult_t pcclExtDiLoCoOuterStep(const pcclExtOuterStepDescriptor_t *d, const pcclComm_t *c, pcclReduceInfo_t *info_out) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(c && d, pcclInvalidArgument);
    PCCL_REQUIRE(c->client, pcclInvalidUsage);
    PCCL_REQUIRE(d->params && d->inner && d->momentum && d->count > 0, pcclInvalidArgument);
    // all three tensors live in the same place
    const auto where = device::classify_pointer(d->params);
    for (const void *q : {static_cast<const void *>(d->inner), static_cast<const void *>(d->momentum)}) {
        const auto w = device::classify_pointer(q);
        PCCL_REQUIRE(w.is_device == where.is_device && w.device == where.device, pcclInvalidArgument);
    }
    PCCL_REQUIRE(c->client->local_world() >= 2, pcclTooFewPeers);
    device::OuterStepArgs step;
    step.params = static_cast<float *>(d->params);
    step.inner = static_cast<float *>(d->inner);
    step.momentum = static_cast<float *>(d->momentum);
    step.lr = d->lr;
    step.mu = d->mu;
    step.first = d->first_step != 0;
    client::ReduceRequest req;
    req.count = d->count;
    req.dtype = wire::DType::F32;
    req.op = wire::RedOp::Avg;
    req.tag = d->tag;
    req.outer = &step; // outlives the call: the operation runs synchronously on this thread
    client::ReduceOutcome out;
    switch (c->client->all_reduce_sync(req, out)) {
    case client::Client::Launch::Ok: break;
    case client::Client::Launch::DuplicateTag: return pcclInvalidArgument;
    case client::Client::Launch::NotConnected: return pcclInvalidUsage;
    }
    if (info_out) {
        info_out->local_world_size = out.world;
        info_out->tx_bytes = out.tx_bytes;
        info_out->rx_bytes = out.rx_bytes;
    }
    return out.ok ? pcclSuccess : pcclRankConnectionLost;
}

// Many small reductions with the same element type and operator (the per-parameter pattern of the
// reference's DiLoCo example, sync_diloco.py:500-505) are fused: gathered into one buffer, reduced
// by ONE operation, scattered back.  Every peer passes the same descriptor list, so every peer
// takes the same decision.  Returns pcclInternalError when the list does not qualify (the caller
// then runs the general scheduler).  PCCL_FUSE_MULTIPLE=0 switches this off.
static pcclResult_t try_fused_multiple(const pcclReduceOpDescriptor_t *descriptors, size_t count, const pcclComm_t *c,
                                       pcclReduceInfo_t *info_out) {
    static const bool enabled = [] {
        const char *v = std::getenv("PCCL_FUSE_MULTIPLE");
        return !(v && v[0] == '0');
    }();
    if (!enabled || count < 2) return pcclInternalError;
    const pcclReduceDescriptor_t &first = descriptors[0].descriptor;
    wire::DType dtype;
    if (!to_wire_dtype(first.src_descriptor.datatype, dtype) || dtype == wire::DType::F8E4M3) return pcclInternalError;
    const size_t esz = wire::dtype_size(dtype);
    size_t total_bytes = 0;
    bool any_device = false, any_host = false;
    for (size_t i = 0; i < count; ++i) {
        const auto &d = descriptors[i];
        if (!d.sendbuf || !d.recvbuf || d.descriptor.count == 0) return pcclInvalidArgument;
        if (d.descriptor.src_descriptor.datatype != first.src_descriptor.datatype || d.descriptor.op != first.op ||
            d.descriptor.quantization_options.algorithm != pcclQuantNone)
            return pcclInternalError;
        (device::classify_pointer(d.sendbuf).is_device ? any_device : any_host) = true;
        (device::classify_pointer(d.recvbuf).is_device ? any_device : any_host) = true;
        total_bytes += (d.descriptor.count * esz + 15) / 16 * 16;
    }
    if (any_device == any_host || total_bytes > (size_t(1) << 30)) return pcclInternalError;
    device::Plane &plane = c->client->plane();
    std::vector<uint8_t> host_fused;
    uint8_t *fused = nullptr;
    std::vector<device::Plane::Part> in_parts, out_parts;
    for (size_t i = 0; i < count; ++i) {
        in_parts.push_back({descriptors[i].sendbuf, descriptors[i].descriptor.count * esz});
        out_parts.push_back({descriptors[i].recvbuf, descriptors[i].descriptor.count * esz});
    }
    std::unique_lock<std::mutex> scratch_guard;
    if (any_device) {
        scratch_guard = plane.fuse_lock(); // one buffer per plane: held until the results are scattered back
        fused = static_cast<uint8_t *>(plane.fuse_buffer(total_bytes));
        if (!fused || !plane.gather(in_parts, fused, true)) return pcclInternalError;
    } else {
        auto *comm = const_cast<pcclComm_t *>(c);
        scratch_guard = std::unique_lock<std::mutex>(comm->fuse_mu, std::try_to_lock);
        if (scratch_guard.owns_lock()) {
            if (comm->fuse_bytes < total_bytes) {
                comm->fuse_buf.reset(new uint8_t[total_bytes]); // not value-initialised: every byte that matters is copied in below
                comm->fuse_bytes = total_bytes;
            }
            fused = comm->fuse_buf.get();
        } else { // a second list in flight on the same communicator: its own buffer
            host_fused.resize(total_bytes);
            fused = host_fused.data();
        }
        size_t off = 0;
        for (const auto &p : in_parts) {
            std::memcpy(fused + off, p.ptr, p.bytes);
            const size_t padded = (p.bytes + 15) / 16 * 16;
            std::memset(fused + off + p.bytes, 0, padded - p.bytes);
            off += padded;
        }
    }
    // padding between the parts is reduced too; its contents are irrelevant but must be defined
    pcclReduceDescriptor_t desc = first;
    desc.count = total_bytes / esz;
    client::ReduceRequest req;
    if (const pcclResult_t rc = build_request(fused, fused, &desc, req); rc != pcclSuccess) return rc;
    req.input_is_scratch = !any_device; // a failed attempt is followed by a fresh gather below, no rollback copy needed
    client::ReduceOutcome out;
    for (;;) {
        if (c->client->local_world() < 2) {
            if (info_out) *info_out = {static_cast<uint32_t>(c->client->local_world()), 0, 0};
            return pcclTooFewPeers;
        }
        const auto launch = c->client->all_reduce_sync(req, out);
        if (launch == client::Client::Launch::NotConnected) return pcclInvalidUsage;
        if (launch == client::Client::Launch::DuplicateTag) return pcclInvalidArgument;
        if (out.ok) break;
        // The fused buffer may live in symmetric memory, where a failed in-place attempt can leave partly reduced
        // values behind: bring the inputs in again before the next attempt.
        if (any_device) {
            if (!plane.gather(in_parts, fused, true)) return pcclInternalError;
        } else {
            size_t off = 0;
            for (const auto &p : in_parts) {
                std::memcpy(fused + off, p.ptr, p.bytes);
                off += (p.bytes + 15) / 16 * 16;
            }
        }
    }
    if (any_device) {
        if (!plane.gather(out_parts, fused, false)) return pcclInternalError;
    } else {
        size_t off = 0;
        for (const auto &p : out_parts) {
            std::memcpy(p.ptr, fused + off, p.bytes);
            off += (p.bytes + 15) / 16 * 16;
        }
    }
    if (info_out) *info_out = {out.world, out.tx_bytes, out.rx_bytes};
    return pcclSuccess;
}

// Scheduler: keep up to max_in_flight reductions running; when any fails, drain everything
// that is in flight (they share the broken ring) and start the unfinished ones again on the
// new ring.  Ends when all are done or the group shrank below two peers.
pcclResult_t pcclAllReduceMultipleWithRetry(const pcclReduceOpDescriptor_t *descriptors, size_t count,
                                            const pcclComm_t *c, pcclReduceInfo_t *info_out, int max_in_flight) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(descriptors && count > 0 && c && max_in_flight > 0, pcclInvalidArgument);
    PCCL_REQUIRE(c->client, pcclInvalidUsage);
    if (const pcclResult_t fused = try_fused_multiple(descriptors, count, c, info_out); fused != pcclInternalError) return fused;
    std::vector<bool> done(count, false);
    size_t n_done = 0;
    uint64_t tx = 0, rx = 0;
    uint32_t world = 0;
    while (n_done < count) {
        std::vector<std::pair<size_t, pcclAsyncReduceOp_t>> flight;
        size_t next = 0;
        bool failed = false;
        auto launch_more = [&]() -> pcclResult_t {
            while (!failed && flight.size() < static_cast<size_t>(max_in_flight)) {
                while (next < count && done[next]) ++next;
                if (next >= count) break;
                pcclAsyncReduceOp_t h{};
                const pcclResult_t rc = pcclAllReduceAsync(descriptors[next].sendbuf, descriptors[next].recvbuf,
                                                           &descriptors[next].descriptor, c, &h);
                if (rc != pcclSuccess) return rc;
                flight.emplace_back(next, h);
                ++next;
            }
            return pcclSuccess;
        };
        pcclResult_t rc = launch_more();
        if (rc != pcclSuccess && flight.empty()) return rc;
        pcclResult_t hard_error = rc;
        while (!flight.empty()) {
            const auto [idx, h] = flight.front();
            flight.erase(flight.begin());
            pcclReduceInfo_t info{};
            const pcclResult_t arc = pcclAwaitAsyncReduce(&h, &info);
            world = info.local_world_size;
            if (arc == pcclSuccess) {
                done[idx] = true;
                ++n_done;
                tx += info.tx_bytes;
                rx += info.rx_bytes;
                if (hard_error == pcclSuccess) {
                    rc = launch_more();
                    if (rc != pcclSuccess) hard_error = rc;
                }
            } else {
                failed = true; // stop launching, drain the rest
            }
        }
        if (hard_error != pcclSuccess && hard_error != pcclTooFewPeers) return hard_error;
        if (n_done < count && c->client->local_world() < 2) {
            if (info_out) *info_out = {static_cast<uint32_t>(c->client->local_world()), tx, rx};
            return pcclTooFewPeers;
        }
    }
    if (info_out) *info_out = {world, tx, rx};
    return pcclSuccess;
}

pcclResult_t pcclSynchronizeSharedState(const pcclComm_t *c, pcclSharedState_t *shared_state,
                                        pcclSharedStateSyncStrategy_t strategy, pcclSharedStateSyncInfo_t *info_out) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(c && shared_state, pcclInvalidArgument);
    PCCL_REQUIRE(c->client, pcclInvalidUsage);
    PCCL_REQUIRE(!c->client->any_reduce_running(), pcclPendingAsyncOps);
    PCCL_REQUIRE(static_cast<int>(strategy) >= 0 && static_cast<int>(strategy) <= 2, pcclInvalidArgument);
    client::SharedStateArgs args;
    args.revision = shared_state->revision;
    args.strategy = static_cast<wire::SyncStrategy>(strategy);
    std::unordered_set<std::string> names;
    for (size_t i = 0; i < shared_state->count; ++i) {
        const pcclTensorInfo_t &t = shared_state->infos[i];
        PCCL_REQUIRE(t.name && t.data, pcclInvalidArgument);
        client::SharedEntry e;
        e.key = t.name;
        PCCL_REQUIRE(names.insert(e.key).second, pcclInvalidArgument); // duplicate key
        e.data = t.data;
        e.count = t.count;
        PCCL_REQUIRE(to_wire_dtype(t.datatype, e.dtype), pcclInvalidArgument);
        e.allow_content_inequality = t.allow_content_inequality;
        const auto where = device::classify_pointer(t.data);
        e.on_device = t.device_type == pcclDeviceCuda || where.is_device;
        args.entries.push_back(std::move(e));
    }
    uint64_t tx = 0, rx = 0;
    if (!c->client->sync_shared_state(args, tx, rx)) return pcclInvalidUsage;
    shared_state->revision = args.revision;
    if (info_out) {
        info_out->tx_bytes = tx;
        info_out->rx_bytes = rx;
    }
    return pcclSuccess;
}

pcclResult_t pcclCreateMaster(ccoip_socket_address_t listen_address, pcclMasterInstance_t **out) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(out, pcclInvalidArgument);
    auto *m = new pcclMasterInstanceState_t;
    m->master = std::make_unique<master::Master>(listen_address);
    *out = m;
    return pcclSuccess;
}

pcclResult_t pcclRunMaster(pcclMasterInstance_t *m) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(m, pcclInvalidArgument);
    PCCL_REQUIRE(!m->running, pcclInvalidArgument);
    if (!m->master->run()) return pcclInternalError;
    m->running = true;
    return pcclSuccess;
}

pcclResult_t pcclInterruptMaster(pcclMasterInstance_t *m) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(m, pcclInvalidArgument);
    if (!m->master->interrupt()) return pcclInternalError;
    m->interrupted = true;
    return pcclSuccess;
}

pcclResult_t pcclMasterAwaitTermination(pcclMasterInstance_t *m) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(m, pcclInvalidArgument);
    PCCL_REQUIRE(m->running, pcclInvalidArgument);
    if (!m->master->join()) return pcclInvalidArgument;
    m->awaited = true;
    return pcclSuccess;
}

pcclResult_t pcclDestroyMaster(pcclMasterInstance_t *m) {
    PCCL_REQUIRE_INIT();
    PCCL_REQUIRE(m, pcclInvalidArgument);
    delete m;
    return pcclSuccess;
}

pcclResult_t pcclExtMasterGetPort(const pcclMasterInstance_t *m, uint16_t *port_out) {
    PCCL_REQUIRE(m && port_out, pcclInvalidArgument);
    *port_out = m->master->port();
    return pcclSuccess;
}

pcclResult_t pcclExtGetLastReduceStats(const pcclComm_t *c, uint64_t tag, pcclExtReduceStats_t *stats_out) {
    PCCL_REQUIRE(c && stats_out, pcclInvalidArgument);
    PCCL_REQUIRE(c->client, pcclInvalidUsage);
    return c->client->last_stats(tag, *stats_out) ? pcclSuccess : pcclInvalidArgument;
}

pcclResult_t pcclExtHashBuffer(const void *data, size_t n_bytes, pcclDeviceType_t device_type, int hash_type,
                               uint64_t *hash_out) {
    PCCL_REQUIRE(hash_out && (data || n_bytes == 0), pcclInvalidArgument);
    PCCL_REQUIRE(hash_type >= 0 && hash_type <= 2, pcclInvalidArgument);
    const bool on_device = device_type == pcclDeviceCuda || device::classify_pointer(data).is_device;
    if (on_device) {
        static std::unique_ptr<device::Plane> plane = device::Plane::create();
        return plane->hash_device(data, n_bytes, static_cast<wire::HashType>(hash_type), *hash_out) ? pcclSuccess
                                                                                                     : pcclInternalError;
    }
    *hash_out = hash_type == 0 ? cpu::simplehash(data, n_bytes) : hash_type == 1 ? cpu::crc32c(data, n_bytes) : cpu::mset64(data, n_bytes);
    return pcclSuccess;
}

pcclResult_t pcclExtSymmetricAlloc(pcclComm_t *c, size_t bytes, void **ptr_out) {
    PCCL_REQUIRE(c && ptr_out && bytes > 0, pcclInvalidArgument);
    PCCL_REQUIRE(c->client, pcclInvalidUsage);
    *ptr_out = c->client->plane().symmetric_alloc(bytes);
    return *ptr_out ? pcclSuccess : pcclInternalError;
}

pcclResult_t pcclExtSymmetricFree(pcclComm_t *c, void *ptr) {
    PCCL_REQUIRE(c && ptr, pcclInvalidArgument);
    PCCL_REQUIRE(c->client, pcclInvalidUsage);
    return c->client->plane().symmetric_free(ptr) ? pcclSuccess : pcclInvalidArgument;
}

} // extern "C"
