// The CCoIP master: a single-threaded micro-consensus state machine.  Every client message is a
// vote; when all peers that must vote have voted the master broadcasts the decision.
//
// Behavioural specification: reference ccoip/src/cpp/ccoip_master_handler.cpp:64-1504 and
// ccoip/src/cpp/ccoip_master_state.cpp:9-1889 (state enums: ccoip/internal/ccoip_master_state.hpp:15-213),
// plus the contracts distilled in SURVEY.md appendix A/B.  The implementation is organised
// differently: one `Master` object owns the server socket and all state; per-peer-group data
// lives in a `Group` record; derived shared-state data (mask, dirty keys) is recomputed from
// the votes on every evaluation, so re-evaluating after a kick or a disconnect is idempotent.
//
// Deliberate deviations from the reference (all documented in DESIGN.md):
//   * reachable-ring construction uses the correct predicate (an edge exists when NEITHER side
//     reported the other unreachable); the reference has it inverted (ccoip_master_state.cpp:1705-1710).
//   * a disconnect during topology optimisation re-evaluates the optimisation votes, so the
//     remaining peers do not hang.
//   * a shared-state entry whose element count differs from the elected mask is a key-set
//     mismatch (the reference ignores the count and would overrun the receiver's buffer).
#pragma once
#include <atomic>
#include <future>
#include <map>
#include <memory>
#include <mutex>
#include <optional>
#include <set>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "../net/epoll_server.hpp"
#include "../topo/bandwidth_store.hpp"
#include "../util/guard.hpp"
#include "../util/log.hpp"
#include "../util/threadpool.hpp"
#include "../wire/packets.hpp"

namespace pccl::master {

enum class Phase { Registered, Accepted };

enum class State {
    Idle,
    VoteAcceptNewPeers,
    VoteNoNewPeersEstablishP2P,
    ConnectingToPeers,
    ConnectingToPeersFailed,
    WaitingForOtherPeers,
    VoteOptimizeTopology,
    OptimizeTopology,
    VoteCompleteTopologyOptimization,
    OptimizeTopologyFailed,
    VoteSyncSharedState,
    DistributeSharedState,
    RequestSharedState,
    VoteCompleteSharedStateSync,
    CollectiveCommunicationsRunning
};

enum class CollectiveState { VoteInitiate, Perform, VoteComplete };

enum class RevisionStatus { Match, Outdated };
enum class ContentStatus { Match, ContentMismatch, KeySetMismatch };

struct Client {
    Uuid uuid;
    ccoip_socket_address_t link{}; // source address of the master connection
    Phase phase = Phase::Registered;
    State state = State::Idle;
    uint32_t group = 0;
    ccoip_socket_address_t p2p_addr{}, ss_addr{}, bm_addr{};
    bool voted_peers_pending = false;
    std::map<uint64_t, CollectiveState> collectives;
    std::map<uint64_t, uint8_t> device_capable; // per tag, between vote and commence (capability bits)
    std::map<uint64_t, uint8_t> queued_initiate; // next vote on a tag whose current operation is completing
};

class Master {
public:
    explicit Master(const ccoip_socket_address_t &listen_address);
    ~Master();

    bool run();       // bind + start the loop thread (non-blocking)
    bool interrupt(); // idempotent
    bool join();
    uint16_t port() const { return server_.port(); }

private:
    struct SyncVote {
        Uuid uuid;
        wire::SyncStrategy strategy;
        RevisionStatus rev_status;
        std::vector<wire::SharedStateEntry> entries;
        uint64_t revision = 0;
    };
    struct Group {
        // shared-state sync
        std::vector<SyncVote> sync_votes;       // in arrival order
        uint64_t next_revision = 0;
        bool revision_initialised = false;
        std::unordered_set<Uuid> sync_complete_votes;
        // collectives
        std::unordered_map<uint64_t, bool> aborted;
        // tags whose collectives the group currently runs under a lease (no votes), see eval_collective_initiate
        std::set<uint64_t> leased_tags;
        // topology
        topo::BandwidthStore bandwidth;
        std::vector<Uuid> ring;
        bool ring_optimal = false;
        bool optimised_once = false;
        std::future<void> moonshot;
    };
    enum class Verdict { NotReady, Done, Failed };

    // ---- socket callbacks (loop thread)
    void on_frame(const ccoip_socket_address_t &from, std::span<const uint8_t> frame);
    void on_disconnect(const ccoip_socket_address_t &from);

    // ---- packet handlers
    void handle_registration(const ccoip_socket_address_t &from, const wire::C2MSessionRegistration &p);
    void handle_establish_p2p(Client &c, const wire::C2MEstablishP2P &p);
    void handle_p2p_established(Client &c, const wire::C2MP2PEstablished &p);
    void handle_peers_pending(Client &c);
    void handle_optimize_topology(Client &c);
    void handle_bandwidth_report(Client &c, const wire::C2MReportBandwidth &p);
    void handle_optimize_work_complete(Client &c);
    void handle_sync_shared_state(Client &c, const wire::C2MSyncSharedState &p);
    void handle_dist_complete(Client &c);
    void handle_collective_initiate(Client &c, const wire::C2MCollectiveInitiate &p);
    void handle_collective_complete(Client &c, const wire::C2MCollectiveComplete &p);

    // ---- consensus evaluation
    void eval_establish_votes();
    void begin_establishment(bool accept_new_peers);
    void eval_p2p_established();
    void eval_peers_pending();
    void eval_optimize_votes();
    void eval_optimize_complete();
    Verdict eval_sync_votes(uint32_t group);
    void eval_sync_complete(uint32_t group);
    void eval_collective_initiate(uint32_t group, uint64_t tag);
    void eval_collective_complete(uint32_t group, uint64_t tag);
    void broadcast_abort(uint32_t group, uint64_t tag, bool aborted);
    bool mark_aborted(uint32_t group, uint64_t tag);
    void revoke_lease(uint32_t group, uint64_t tag, bool all, bool peer_lost);

    // ---- helpers
    Client *find(const ccoip_socket_address_t &addr);
    Client *find(const Uuid &uuid);
    void kick(const ccoip_socket_address_t &addr, const char *why);
    template <typename P> void send(const Client &c, const P &pkt) {
        if (!server_.send_packet(c.link, pkt)) PLOG(DEBUG) << "send to " << net::to_string(c.link) << " failed";
    }
    std::vector<Client *> ordered_clients();
    void send_connection_info(Client &c, bool include_registered);
    std::vector<Uuid> ring_for(uint32_t group, bool include_registered);
    std::vector<Uuid> basic_ring(uint32_t group);
    std::optional<std::vector<Uuid>> reachable_ring(uint32_t group);
    uint64_t local_world(uint32_t group, bool include_registered) const;
    uint64_t num_groups(bool include_registered) const;
    uint64_t largest_group(bool include_registered) const;
    void promote(Client &c);
    void run_ring_optimisation(uint32_t group);
    std::vector<topo::Edge> missing_edges(const Client &c);

    net::EpollServer server_;
    std::atomic<bool> running_{false}, interrupted_{false};
    ThreadGuard state_owner_; // everything below belongs to the server's event thread

    std::unordered_map<Uuid, Client> clients_;
    std::unordered_map<net::AddrKey, Uuid> by_addr_;
    std::map<uint32_t, Group> groups_;
    std::unordered_map<Uuid, std::vector<Uuid>> last_p2p_info_;
    std::unordered_map<Uuid, std::unordered_set<Uuid>> unreachable_;
    uint64_t next_seq_nr_ = 0;
    bool peer_dropped_ = false;

    std::mutex topo_mu_; // ring / ring_optimal are also touched by the optimiser thread
    ThreadPool topo_pool_{2, 16};
};

} // namespace pccl::master
