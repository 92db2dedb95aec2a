// A peer of a CCoIP run: master link, P2P links to the ring neighbours, shared-state server,
// bandwidth-benchmark server, asynchronous all-reduce engine and the device data plane.
//
// Behavioural specification: reference ccoip/src/cpp/ccoip_client_handler.cpp (connect :46-280,
// P2P establishment :282-314,831-1002, shared state :344-638,1022-1164, topology optimisation :640-736,
// all-reduce :1187-1396) and ccoip/src/cpp/ccoip_client_state.cpp:184-232.  The structure is new:
// one class, condition-variable based waiting, posted zero-copy sends, a pluggable device plane.
#pragma once
#include <atomic>
#include <condition_variable>
#include <deque>
#include <future>
#include <map>
#include <memory>
#include <mutex>
#include <optional>
#include <set>
#include <shared_mutex>
#include <unordered_map>
#include <vector>

#include "../device/plane.hpp"
#include "../net/blocking_socket.hpp"
#include "../net/mux_socket.hpp"
#include "../net/queued_socket.hpp"
#include "../util/threadpool.hpp"
#include "../util/uuid.hpp"
#include "../wire/packets.hpp"
#include "pccl.h"

namespace pccl::client {

struct ClientConfig {
    ccoip_socket_address_t master{};
    uint32_t peer_group = 0;
    uint32_t pool_size = 1;
    bool explicit_addresses = false;
    ccoip_socket_address_t adv_p2p{}, adv_ss{}, adv_bm{};
    uint16_t p2p_port = CCOIP_PROTOCOL_PORT_P2P, ss_port = CCOIP_PROTOCOL_PORT_SHARED_STATE,
             bm_port = CCOIP_PROTOCOL_PORT_BANDWIDTH_BENCHMARK;
    // Two-level all-reduce for groups that span hosts (see Client::build_hierarchy).  host_id 0 = derive it (PCCL_HOST_ID, else
    // host name + boot id); hierarchy -1 = automatic (PCCL_HIERARCHY, else on when the master is not on loopback or a host id
    // was given), 0 = off, 1 = on.  Tests set these per client to emulate several hosts in one process.
    uint64_t host_id = 0;
    int hierarchy = -1;
};

struct SharedEntry {
    std::string key;
    void *data = nullptr;
    uint64_t count = 0;
    wire::DType dtype = wire::DType::U8;
    bool on_device = false;
    bool allow_content_inequality = false;
    size_t bytes() const { return count * wire::dtype_size(dtype); }
};

struct SharedStateArgs {
    uint64_t revision = 0;
    std::vector<SharedEntry> entries;
    wire::SyncStrategy strategy = wire::SyncStrategy::EnforcePopular;
};

struct ReduceRequest {
    const void *send = nullptr;
    void *recv = nullptr;
    uint64_t count = 0;
    wire::DType dtype = wire::DType::F32;
    wire::RedOp op = wire::RedOp::Sum;
    wire::QuantAlgo quant = wire::QuantAlgo::None;
    wire::DType quant_dtype = wire::DType::U8;
    uint64_t tag = 0;
    // DiLoCo outer step: when set, send/recv are ignored; the reduction averages params - inner (fp32) and
    // its result drives the Nesterov update of params/momentum/inner (see device::OuterStepArgs)
    const device::OuterStepArgs *outer = nullptr;
    // stream-ordered form (device buffers): the operation is ordered after the work already queued on this CUDA stream
    // and the stream waits for its result; with `enqueue` the call returns after the launch when no vote is needed
    void *user_stream = nullptr;
    bool enqueue = false;
    bool want_hash = false; // also produce the mset64 content hash(es) of the result (device plane, one launch)
    // the operand is a scratch copy the caller can make again: an in-place host reduction then keeps no backup to roll back to
    bool input_is_scratch = false;
};

struct ReduceOutcome {
    bool ok = false;
    uint32_t world = 0;
    uint64_t tx_bytes = 0, rx_bytes = 0;
    pcclExtReduceStats_t stats{};
    bool pending = false;  // enqueued on the device, verdict not taken yet (see Client::wait_stream_ops)
    uint64_t ticket = 0;
    bool has_hash = false;
    uint64_t hash[2] = {0, 0};
};

class Client {
public:
    explicit Client(const ClientConfig &cfg);
    ~Client();

    bool connect();
    bool interrupt();
    bool join();
    bool is_connected() const { return accepted_.load(); }

    // joint operations (one caller thread at a time)
    bool update_topology();                 // accept pending peers + rewire
    bool are_peers_pending(bool &pending);
    bool optimize_topology();
    bool sync_shared_state(SharedStateArgs &state, uint64_t &tx_bytes, uint64_t &rx_bytes);

    // asynchronous reductions (any thread)
    enum class Launch { Ok, DuplicateTag, NotConnected };
    Launch all_reduce_async(const ReduceRequest &req);
    bool await_reduce(uint64_t tag, ReduceOutcome &out);
    // Blocking variant that runs on the calling thread (no hand-over to the worker pool).
    Launch all_reduce_sync(const ReduceRequest &req, ReduceOutcome &out);
    // Stream-ordered reduction: like all_reduce_sync, but ordered after req.user_stream; when the operation runs under a
    // lease the call returns right after the launch and the verdict is taken by the next operation on the same tag or
    // by wait_stream_ops().  Several may be in flight, also on one tag (the stream orders them).
    Launch all_reduce_on_stream(const ReduceRequest &req, ReduceOutcome &out);
    // Takes the verdict of the enqueued operation on `tag` (all tags when `all`).  False when one of them failed.
    bool wait_stream_ops(uint64_t tag, bool all, ReduceOutcome &last);
    bool verify_received(const std::vector<std::pair<const SharedEntry *, uint64_t>> &to_verify, wire::HashType hash_type);
    bool any_reduce_running();
    bool last_stats(uint64_t tag, pcclExtReduceStats_t &out);

    // Content hashes the collective kernels produced while writing a tensor (mset64).  With trust switched on, a shared-state
    // synchronisation that uses the mset64 hash type takes a tensor's hash from here instead of launching a hash kernel --
    // valid as long as nothing but this library wrote the tensor since (the caller's promise when switching it on).
    void trust_content_hashes(bool on) { trust_hashes_ = on; }
    bool trusts_content_hashes() const { return trust_hashes_; }
    void remember_hash(const void *ptr, size_t bytes, uint64_t value);
    void forget_hashes(const void *ptr, size_t bytes); // anything overlapping [ptr, ptr+bytes)
    bool recall_hash(const void *ptr, size_t bytes, uint64_t &value);
    uint64_t hash_kernels_skipped() const { return hash_skips_; }

    uint64_t global_world() const { return global_world_; }
    // ports this peer listens on (after connect()): p2p, shared state, bandwidth benchmark
    uint16_t p2p_port() const { return p2p_listener_.port(); }
    uint16_t shared_state_port() const { return ss_listener_.port(); }
    uint16_t benchmark_port() const { return bm_listener_.port(); }
    uint64_t local_world() const { return local_world_; }
    uint64_t num_groups() const { return num_groups_; }
    uint64_t largest_group() const { return largest_group_; }
    // (hosts, peers per host) of the two-level layout in use, (0, 0) when the ring is flat
    std::pair<uint32_t, uint32_t> hierarchy_shape() {
        std::shared_lock<std::shared_mutex> lock(ring_mu_);
        return hier_.ready ? std::make_pair(hier_.B, hier_.L) : std::make_pair(0u, 0u);
    }
    const Uuid &uuid() const { return uuid_; }
    device::Plane &plane() { return *plane_; }

private:
    enum class Establish { Success, Retry, Failed };
    using Pool = std::vector<std::shared_ptr<net::MuxSocket>>;

    bool request_and_establish(bool accept_new_peers);
    Establish establish_round();
    bool dial_peer(const wire::PeerEndpoint &peer);
    void on_p2p_accept(std::unique_ptr<net::BlockingSocket> sock);
    void on_shared_state_accept(std::unique_ptr<net::BlockingSocket> sock);
    void on_benchmark_accept(std::unique_ptr<net::BlockingSocket> sock);
    ReduceOutcome run_reduce(const ReduceRequest &req, uint64_t revision_at_launch);
    ReduceOutcome run_reduce_impl(const ReduceRequest &req, uint64_t revision_at_launch);
    bool conclude(uint64_t tag, uint64_t revision_at_launch, ReduceOutcome &out);
    struct StreamOp;
    bool finish_stream_op(uint64_t tag, const StreamOp &so, ReduceOutcome &out);
    // Completion handshakes that were left pending after a device-plane reduction returned early.
    struct Deferred {
        bool abort_seen = false, aborted = false;
    };
    bool finish_deferred(uint64_t tag, Deferred &d);
    bool drain_deferred();
    Pool pick_rx_pool(const Uuid &peer, int timeout_ms);
    uint8_t *staging_for(uint64_t tag, size_t bytes);
    bool run_benchmark(const wire::BenchmarkRequest &req, double &mbits, bool &busy);

    ClientConfig cfg_;
    net::QueuedSocket master_;
    net::Listener p2p_listener_, ss_listener_, bm_listener_;
    std::unique_ptr<device::Plane> plane_;

    Uuid uuid_;
    std::atomic<bool> accepted_{false}, interrupted_{false}, started_{false};
    std::atomic<uint64_t> global_world_{0}, local_world_{0}, num_groups_{0}, largest_group_{0};
    std::atomic<uint64_t> connection_revision_{0};

    std::mutex joint_mu_; // serialises joint (phase changing) calls and re-establishment

    std::shared_mutex ring_mu_;
    std::vector<Uuid> ring_;
    // Two-level layout of the ring (guarded by ring_mu_): L peers on each of B hosts; the local ring links the peers of this
    // host, the cross ring the peers of equal local rank on all hosts.
    struct Hierarchy {
        bool ready = false;
        uint32_t L = 0, B = 0, local_rank = 0, box = 0;
        Uuid local_next, local_prev, cross_next, cross_prev;
    };
    Hierarchy hier_;
    std::vector<wire::PeerEndpoint> last_neighbours_; // endpoints of the ring neighbours, from the master's last full connection info
    uint64_t host_id_ = 0;
    bool hierarchy_enabled_ = false;
    void build_hierarchy(const std::vector<Uuid> &ring);

    std::mutex tx_mu_;
    std::unordered_map<Uuid, Pool> tx_pools_;
    std::mutex rx_mu_;
    std::condition_variable rx_cv_;
    std::unordered_map<Uuid, Pool> rx_pools_;

    struct Op {
        std::future<ReduceOutcome> future;
        uint64_t revision_at_launch = 0;
        // leased device operations are launched on the calling thread (no hand-over to a worker, whose wake-up costs more
        // than the whole operation); await_reduce() then only collects the verdict
        bool launched_inline = false;
        uint64_t ticket = 0;
        ReduceOutcome partial;
    };
    // page-locked host buffers for device tensors that travel over the TCP plane, one per tag
    struct StageBuf {
        uint8_t *ptr = nullptr;
        size_t bytes = 0;
    };
    std::mutex stage_mu_;
    std::map<uint64_t, StageBuf> stage_cache_;
    // rollback copies of in-place host operands, one per tag, kept between calls: a training loop reduces the same buffers
    // every step, and a fresh vector of that size costs more in page faults than the copy itself
    std::map<uint64_t, std::vector<uint8_t>> backup_cache_;
    uint64_t backup_cache_bytes_ = 0;
    std::mutex deferred_mu_;
    std::map<uint64_t, Deferred> deferred_;
    std::atomic<bool> ring_stale_{false};       // a peer vanished after an operation had finished here: rewire before the next one
    std::atomic<bool> master_pipelines_{false}; // the master queues an Initiate behind a pending Complete
    // Leases: tags whose collectives run without a vote at the master (granted by the Commence of an earlier operation on
    // that tag, see run_reduce), and how many leased operations each tag has seen on this ring (the operation's name in
    // the same-box sequencer).  Everything here is reset when the ring changes.
    std::mutex lease_mu_;
    std::set<uint64_t> lease_tags_;
    std::map<uint64_t, uint64_t> lease_counts_;
    std::atomic<bool> lease_mc_all_{false}; // every peer of the ring holds the multicast view
    bool lease_enabled_ = true;             // PCCL_LEASE=0 keeps every operation a vote
    struct StreamOp {
        uint64_t ticket = 0, revision_at_launch = 0;
        ReduceOutcome partial;
    };
    std::deque<std::pair<uint64_t, StreamOp>> stream_ops_; // (tag, operation) in launch order; guarded by ops_mu_
    std::mutex ops_mu_;
    std::map<uint64_t, Op> ops_;
    std::map<uint64_t, pcclExtReduceStats_t> stats_;
    std::unique_ptr<ThreadPool> workers_;

    struct KnownHash {
        size_t bytes;
        uint64_t value;
    };
    std::mutex hash_mu_;
    std::map<uintptr_t, KnownHash> known_hashes_;
    std::atomic<bool> trust_hashes_{false};
    std::atomic<uint64_t> hash_skips_{0};

    // shared-state serving
    std::mutex ss_mu_;
    bool ss_serving_ = false;
    SharedStateArgs ss_snapshot_;
    std::atomic<uint64_t> ss_tx_bytes_{0};
    std::unique_ptr<ThreadPool> ss_workers_;

    // benchmark serving
    std::mutex bm_mu_;
    Uuid bm_owner_;
    int bm_active_ = 0;
    std::unique_ptr<ThreadPool> bm_workers_;
};

} // namespace pccl::client
