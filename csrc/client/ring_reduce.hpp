// Pipelined ring all-reduce over the multiplexed TCP links: reduce-scatter followed by all-gather,
// W-1 steps each.  Algorithmic contract as in the reference (ccoip/src/cpp/reduce.cpp:51-67 chunking,
// :609-662 reduce-scatter, :673-769 all-gather incl. the "owner adopts D(Q(x))" rule that keeps
// all peers bit-identical under quantisation, :782 AVG finalisation, :572-580 source restore on
// abort).  The implementation differs: sends are zero-copy tickets on the TX thread, received
// frames are reduced straight out of the frame buffer, and waiting is done on condition
// variables with a short timeout that doubles as the abort-poll interval.
#pragma once
#include <functional>
#include <memory>
#include <vector>

#include "../kernels/cpu/quantize.hpp"
#include "../net/mux_socket.hpp"
#include "../wire/packets.hpp"

namespace pccl::client {

struct RingContext {
    uint64_t tag = 0, seq = 0;
    uint32_t rank = 0, world = 1;
    // Connection pools to the ring successor / from the predecessor (same size on both ends).
    // Small transfers use connection seq % pool; large ones are striped over all of them so that
    // several kernel socket queues (and RX/TX threads) work on one operation.
    std::vector<std::shared_ptr<net::MuxSocket>> tx, rx;
    size_t frame_bytes = 2u << 20;
    size_t stripe_threshold = 512u << 10;
    // returns true when the master told us to abort this operation
    std::function<bool()> abort_requested;
};

struct RingResult {
    bool ok = false;
    bool aborted_by_master = false;
    uint64_t tx_bytes = 0, rx_bytes = 0;
};

struct ReduceArgs {
    const void *src = nullptr;
    void *dst = nullptr;
    uint64_t count = 0;
    wire::DType dtype = wire::DType::F32;
    wire::RedOp op = wire::RedOp::Sum;
    wire::QuantAlgo quant = wire::QuantAlgo::None;
    wire::DType quant_dtype = wire::DType::U8;
};

// `backup`, when given, receives a copy of the input if the operation is in place (src == dst).
// On a local failure the input is restored from it before returning; on local success it is left
// filled so the caller can still roll back if the master later reports that another peer
// aborted.  Callers that reduce a private staging copy pass nullptr and skip that pass over memory.
RingResult ring_allreduce(const RingContext &ctx, const ReduceArgs &args, std::vector<uint8_t> *backup);

// Two-level all-reduce for groups that span several hosts (no counterpart in the reference, whose ring is flat:
// ccoip/src/cpp/reduce.cpp:528-784): L peers per host, B hosts.
//   1. reduce-scatter over the LOCAL ring (peers of one host): local rank l ends up with the host's sum of chunk (l+1) mod L
//   2. all-reduce of that chunk over the CROSS ring (the peers with the same local rank, one per host)
//   3. all-gather over the local ring
// Every byte crosses the inter-host links 2(B-1)/B times per local-rank lane instead of travelling around one ring of L*B
// hops, and each inter-host link carries 1/L of the data.  A wire quantisation applies to step 2 only (codes between hosts, plain
// values inside a host).  `local` / `cross` carry the rank, world and
// connection pools of the two rings; tag, seq and abort callback are taken from `local`.
RingResult hierarchical_allreduce(const RingContext &local, const RingContext &cross, const ReduceArgs &args, std::vector<uint8_t> *backup);

// chunk i of `count` elements split over `world` owners: [begin, end)
inline void chunk_bounds(uint64_t count, uint32_t world, uint32_t i, uint64_t &begin, uint64_t &end) {
    const uint64_t base = count / world, rem = count % world;
    begin = i * base + (i < rem ? i : rem);
    end = begin + base + (i < rem ? 1 : 0);
}

} // namespace pccl::client
