#include "master.hpp"

#include <algorithm>
#include <functional>

#include "../topo/optimizer.hpp"
#include "../util/log.hpp"

namespace pccl::master {

using namespace pccl::wire;

Master::Master(const ccoip_socket_address_t &listen_address) : server_(listen_address, false) {
    server_.on_read([this](const ccoip_socket_address_t &from, std::span<const uint8_t> frame) { on_frame(from, frame); });
    server_.on_close([this](const ccoip_socket_address_t &from) { on_disconnect(from); });
}

Master::~Master() {
    interrupt();
    join();
    topo_pool_.shutdown();
}

bool Master::run() {
    if (running_) return false;
    if (!server_.listen()) return false;
    if (!server_.run_async()) return false;
    running_ = true;
    PLOG(INFO) << "master listening on port " << server_.port();
    return true;
}

bool Master::interrupt() {
    if (interrupted_.exchange(true)) return true;
    return server_.interrupt();
}

bool Master::join() {
    if (!running_) return false;
    server_.join();
    return true;
}

// ------------------------------------------------------------------------------------------
// lookup helpers
// ------------------------------------------------------------------------------------------

Client *Master::find(const ccoip_socket_address_t &addr) {
    const auto it = by_addr_.find(net::AddrKey(addr));
    if (it == by_addr_.end()) return nullptr;
    return find(it->second);
}

Client *Master::find(const Uuid &uuid) {
    const auto it = clients_.find(uuid);
    return it == clients_.end() ? nullptr : &it->second;
}

void Master::kick(const ccoip_socket_address_t &addr, const char *why) {
    PLOG(WARN) << "kicking " << net::to_string(addr) << ": " << why;
    server_.close_client(addr);
}

// Deterministic iteration order (by link address) for everything that broadcasts.
std::vector<Client *> Master::ordered_clients() {
    std::vector<Client *> out;
    out.reserve(clients_.size());
    for (auto &[u, c] : clients_) out.push_back(&c);
    std::sort(out.begin(), out.end(),
              [](const Client *a, const Client *b) { return net::AddrKey(a->link) < net::AddrKey(b->link); });
    return out;
}

uint64_t Master::local_world(uint32_t group, bool include_registered) const {
    uint64_t n = 0;
    for (const auto &[u, c] : clients_)
        if (c.group == group && (include_registered || c.phase == Phase::Accepted)) ++n;
    return n;
}

uint64_t Master::num_groups(bool include_registered) const {
    std::set<uint32_t> g;
    for (const auto &[u, c] : clients_)
        if (include_registered || c.phase == Phase::Accepted) g.insert(c.group);
    return g.size();
}

uint64_t Master::largest_group(bool include_registered) const {
    std::map<uint32_t, uint64_t> sizes;
    for (const auto &[u, c] : clients_)
        if (include_registered || c.phase == Phase::Accepted) sizes[c.group]++;
    uint64_t best = 0;
    for (const auto &[g, n] : sizes) best = std::max(best, n);
    return best;
}

void Master::promote(Client &c) {
    if (c.phase == Phase::Accepted) return;
    c.phase = Phase::Accepted;
    groups_[c.group].bandwidth.register_peer(c.uuid);
    PLOG(INFO) << "peer " << c.uuid.str() << " accepted, group " << c.group;
}

// ------------------------------------------------------------------------------------------
// frame dispatch
// ------------------------------------------------------------------------------------------

template <typename P> static bool decode(Reader &r, P &pkt) { return pkt.read(r) && r.ok(); }

void Master::on_frame(const ccoip_socket_address_t &from, std::span<const uint8_t> frame) {
    state_owner_.check("master state machine");
    Reader r(frame.data(), frame.size());
    const PacketId id = r.get<uint16_t>();
    if (!r.ok()) return kick(from, "truncated frame");

    if (id == C2MSessionRegistration::id) {
        C2MSessionRegistration p;
        if (!decode(r, p)) return kick(from, "malformed registration");
        return handle_registration(from, p);
    }
    Client *c = find(from);
    if (!c) return kick(from, "packet from unregistered connection");

    switch (id) {
    case C2MEstablishP2P::id: {
        C2MEstablishP2P p;
        if (!decode(r, p)) return kick(from, "malformed packet");
        return handle_establish_p2p(*c, p);
    }
    case C2MP2PEstablished::id: {
        C2MP2PEstablished p;
        if (!decode(r, p)) return kick(from, "malformed packet");
        return handle_p2p_established(*c, p);
    }
    case C2MCheckPeersPending::id: return handle_peers_pending(*c);
    case C2MOptimizeTopology::id: return handle_optimize_topology(*c);
    case C2MReportBandwidth::id: {
        C2MReportBandwidth p;
        if (!decode(r, p)) return kick(from, "malformed packet");
        return handle_bandwidth_report(*c, p);
    }
    case C2MOptimizeWorkComplete::id: return handle_optimize_work_complete(*c);
    case C2MSyncSharedState::id: {
        C2MSyncSharedState p;
        if (!decode(r, p)) return kick(from, "malformed packet");
        return handle_sync_shared_state(*c, p);
    }
    case C2MDistSharedStateComplete::id: return handle_dist_complete(*c);
    case C2MCollectiveInitiate::id: {
        C2MCollectiveInitiate p;
        if (!decode(r, p)) return kick(from, "malformed packet");
        return handle_collective_initiate(*c, p);
    }
    case C2MCollectiveComplete::id: {
        C2MCollectiveComplete p;
        if (!decode(r, p)) return kick(from, "malformed packet");
        return handle_collective_complete(*c, p);
    }
    default: return kick(from, "unknown packet id");
    }
}

// ------------------------------------------------------------------------------------------
// registration / membership
// ------------------------------------------------------------------------------------------

void Master::handle_registration(const ccoip_socket_address_t &from, const C2MSessionRegistration &p) {
    if (find(from)) return kick(from, "registered twice");

    M2CSessionRegistrationResponse resp;
    // A run is either all-loopback or all-routable: addresses handed out to peers are derived
    // from the source address the master sees.
    bool refuse = false;
    if (!clients_.empty()) {
        bool all_local = true;
        for (const auto &[u, c] : clients_) all_local = all_local && net::is_loopback(c.link);
        if (all_local != net::is_loopback(from)) refuse = true;
    }
    if (refuse) {
        resp.accepted = false;
        server_.send_packet(from, resp);
        return kick(from, "loopback and non-loopback peers cannot be mixed");
    }

    Client c;
    c.uuid = Uuid::random_v4();
    c.link = from;
    c.group = p.peer_group;
    if (p.explicit_addresses) {
        c.p2p_addr = p.adv_p2p;
        c.ss_addr = p.adv_ss;
        c.bm_addr = p.adv_bm;
    } else {
        c.p2p_addr = c.ss_addr = c.bm_addr = from;
        c.p2p_addr.port = p.p2p_port;
        c.ss_addr.port = p.ss_port;
        c.bm_addr.port = p.bm_port;
    }
    const Uuid uuid = c.uuid;
    by_addr_[net::AddrKey(from)] = uuid;
    clients_.emplace(uuid, std::move(c));
    groups_[p.peer_group]; // make sure the group record exists

    resp.accepted = true;
    resp.assigned_uuid = uuid;
    server_.send_packet(from, resp);
    PLOG(INFO) << "peer " << uuid.str() << " registered from " << net::to_string(from) << ", group " << p.peer_group;

    if (clients_.size() == 1) {
        // Nobody to ask: the first peer is accepted on the spot and receives an empty
        // connection-info packet.
        Client &first = clients_.at(uuid);
        promote(first);
        first.state = State::VoteAcceptNewPeers;
    }
    eval_establish_votes();
}

void Master::handle_establish_p2p(Client &c, const C2MEstablishP2P &p) {
    if (c.phase != Phase::Accepted || (c.state != State::Idle && c.state != State::ConnectingToPeersFailed))
        return kick(c.link, "establish-p2p vote in illegal state");
    c.state = p.accept_new_peers ? State::VoteAcceptNewPeers : State::VoteNoNewPeersEstablishP2P;
    eval_establish_votes();
}

// Pending (registered) peers cannot vote but count as agreeing, otherwise the totals never match.
void Master::eval_establish_votes() {
    if (clients_.empty()) return;
    size_t accept = 0, no_accept = 0, registered = 0;
    for (const auto &[u, c] : clients_) {
        if (c.phase == Phase::Registered)
            ++registered;
        else if (c.state == State::VoteAcceptNewPeers)
            ++accept;
        else if (c.state == State::VoteNoNewPeersEstablishP2P)
            ++no_accept;
    }
    if (accept + registered == clients_.size() && (accept > 0 || no_accept == 0))
        begin_establishment(true);
    else if (no_accept + registered == clients_.size() && no_accept > 0)
        begin_establishment(false);
}

void Master::begin_establishment(bool accept_new_peers) {
    peer_dropped_ = false; // a fresh establishment round starts
    // leases belong to a ring: every peer drops its lease state when the round ends (Client::establish_round)
    for (auto &[id, g] : groups_) g.leased_tags.clear();
    for (auto &[u, c] : clients_)
        if (accept_new_peers || c.phase == Phase::Accepted) c.state = State::ConnectingToPeers;
    for (Client *c : ordered_clients())
        if (c->state == State::ConnectingToPeers) send_connection_info(*c, accept_new_peers);
}

std::vector<Uuid> Master::basic_ring(uint32_t group) {
    std::vector<Uuid> ring;
    for (const auto &[u, c] : clients_) {
        // a pending peer that is being wired up right now already belongs to the ring
        if (c.group != group) continue;
        if (c.phase != Phase::Accepted && c.state != State::ConnectingToPeers) continue;
        ring.push_back(u);
    }
    std::sort(ring.begin(), ring.end());
    return ring;
}

std::vector<Uuid> Master::ring_for(uint32_t group, bool include_registered) {
    std::lock_guard<std::mutex> lock(topo_mu_);
    Group &g = groups_[group];
    bool dirty = g.ring.size() != local_world(group, include_registered);
    for (const auto &u : g.ring) dirty = dirty || !clients_.count(u);
    if (dirty) {
        g.ring_optimal = false;
        // prefer an order that avoids pairs known not to reach each other
        auto ring = basic_ring(group);
        bool blocked = false;
        for (size_t i = 0; i < ring.size() && ring.size() > 1; ++i) {
            const Uuid &a = ring[i], &b = ring[(i + 1) % ring.size()];
            if (unreachable_[a].count(b) || unreachable_[b].count(a)) blocked = true;
        }
        if (blocked)
            if (auto alt = reachable_ring(group)) ring = *alt;
        g.ring = std::move(ring);
    }
    return g.ring;
}

// Hamiltonian cycle over pairs that can talk to each other (depth-first with backtracking;
// peer groups are small).
std::optional<std::vector<Uuid>> Master::reachable_ring(uint32_t group) {
    const std::vector<Uuid> nodes = basic_ring(group);
    if (nodes.empty()) return std::nullopt;
    if (nodes.size() == 1) return nodes;
    auto linked = [&](const Uuid &a, const Uuid &b) { return !unreachable_[a].count(b) && !unreachable_[b].count(a); };
    if (nodes.size() == 2) {
        if (linked(nodes[0], nodes[1])) return nodes;
        return std::nullopt;
    }
    std::vector<Uuid> path{nodes[0]};
    std::unordered_set<Uuid> used{nodes[0]};
    std::function<bool()> extend = [&]() -> bool {
        if (path.size() == nodes.size()) return linked(path.back(), path.front());
        for (const auto &cand : nodes) {
            if (used.count(cand) || !linked(path.back(), cand)) continue;
            used.insert(cand);
            path.push_back(cand);
            if (extend()) return true;
            path.pop_back();
            used.erase(cand);
        }
        return false;
    };
    if (extend()) return path;
    return std::nullopt;
}

void Master::send_connection_info(Client &c, bool include_registered) {
    M2CP2PConnectionInfo info;
    info.global_world_size = clients_.size();
    info.local_world_size = local_world(c.group, include_registered);
    info.num_peer_groups = num_groups(include_registered);
    info.largest_group_size = largest_group(include_registered);

    // a peer only talks to its ring neighbours
    std::vector<Uuid> neighbours;
    const auto ring = ring_for(c.group, include_registered);
    const auto me = std::find(ring.begin(), ring.end(), c.uuid);
    if (me != ring.end() && ring.size() > 1) {
        const size_t i = static_cast<size_t>(me - ring.begin());
        const Uuid &prev = ring[(i + ring.size() - 1) % ring.size()];
        const Uuid &next = ring[(i + 1) % ring.size()];
        neighbours.push_back(prev);
        if (next != prev) neighbours.push_back(next);
    }
    auto &before = last_p2p_info_[c.uuid];
    bool changed = before.size() != neighbours.size();
    for (const auto &n : neighbours) changed = changed || std::find(before.begin(), before.end(), n) == before.end();
    before = neighbours;
    info.unchanged = !changed;
    if (changed)
        for (const auto &n : neighbours)
            if (const Client *peer = find(n)) info.peers.push_back({peer->p2p_addr, peer->uuid});
    send(c, info);
}

void Master::handle_p2p_established(Client &c, const C2MP2PEstablished &p) {
    if (c.state != State::ConnectingToPeers) return kick(c.link, "p2p-established report in illegal state");
    if (!p.success) {
        auto &bad = unreachable_[c.uuid];
        for (const auto &peer : p.failed_peers)
            if (find(peer)) bad.insert(peer);
        std::lock_guard<std::mutex> lock(topo_mu_);
        const auto alt = reachable_ring(c.group);
        if (!alt) {
            // cannot be placed into any ring
            server_.close_client(c.link);
            PLOG(WARN) << "no ring exists that avoids the unreachable pairs reported by " << c.uuid.str();
            return;
        }
        Group &g = groups_[c.group];
        g.ring_optimal = false;
        g.ring = *alt;
    }
    c.state = p.success ? State::WaitingForOtherPeers : State::ConnectingToPeersFailed;
    eval_p2p_established();
}

void Master::eval_p2p_established() {
    size_t reported = 0, connecting = 0, failed = 0;
    for (const auto &[u, c] : clients_) {
        if (c.state == State::WaitingForOtherPeers || c.state == State::ConnectingToPeersFailed) ++reported;
        if (c.state == State::ConnectingToPeersFailed) ++failed;
        if (c.state == State::ConnectingToPeers) ++connecting;
    }
    if (reported == 0 || connecting > 0) return;
    const bool failure = failed > 0 || peer_dropped_;
    peer_dropped_ = false;
    for (auto &[u, c] : clients_) {
        if (c.state != State::WaitingForOtherPeers && c.state != State::ConnectingToPeersFailed) continue;
        if (c.state == State::WaitingForOtherPeers) {
            c.state = failure ? State::ConnectingToPeersFailed : State::Idle;
            promote(c); // the newcomer is in, even if the round has to be repeated
        }
    }
    for (Client *c : ordered_clients()) {
        if (c->phase == Phase::Registered) continue; // joined too late for this round
        if (c->state != State::Idle && c->state != State::ConnectingToPeersFailed) continue;
        M2CP2PEstablished out;
        out.success = !failure;
        out.ring_order = ring_for(c->group, false);
        send(*c, out);
    }
}

void Master::handle_peers_pending(Client &c) {
    if (c.phase != Phase::Accepted || c.voted_peers_pending) return kick(c.link, "peers-pending vote in illegal state");
    c.voted_peers_pending = true;
    eval_peers_pending();
}

void Master::eval_peers_pending() {
    bool any_accepted = false, pending = false;
    for (const auto &[u, c] : clients_) {
        if (c.phase == Phase::Registered) {
            pending = true;
            continue;
        }
        any_accepted = true;
        if (!c.voted_peers_pending) return;
    }
    if (!any_accepted) return;
    for (Client *c : ordered_clients()) {
        if (c->phase != Phase::Accepted) continue;
        c->voted_peers_pending = false;
        M2CPeersPendingResponse out;
        out.pending = pending;
        send(*c, out);
    }
}

// ------------------------------------------------------------------------------------------
// topology optimisation
// ------------------------------------------------------------------------------------------

std::vector<topo::Edge> Master::missing_edges(const Client &c) {
    auto edges = groups_[c.group].bandwidth.missing_edges(c.uuid);
    edges.erase(std::remove_if(edges.begin(), edges.end(),
                               [&](const topo::Edge &e) { return unreachable_[e.from].count(e.to) > 0; }),
                edges.end());
    return edges;
}

void Master::handle_optimize_topology(Client &c) {
    if (c.phase != Phase::Accepted || (c.state != State::Idle && c.state != State::OptimizeTopologyFailed))
        return kick(c.link, "optimize-topology vote in illegal state");
    c.state = State::VoteOptimizeTopology;
    eval_optimize_votes();
}

void Master::eval_optimize_votes() {
    size_t votes = 0;
    for (const auto &[u, c] : clients_) {
        if (c.phase == Phase::Registered) continue;
        if (c.state != State::VoteOptimizeTopology) return;
        ++votes;
    }
    if (votes == 0) return;
    for (auto &[u, c] : clients_)
        if (c.phase == Phase::Accepted) c.state = State::OptimizeTopology;
    for (Client *c : ordered_clients()) {
        if (c->phase != Phase::Accepted) continue;
        // a benchmark is run by the sender of the edge, so only edges that start here are sent
        M2COptimizeTopologyResponse out;
        for (const auto &e : missing_edges(*c)) {
            if (e.from != c->uuid) continue;
            const Client *to = find(e.to);
            if (!to || to->phase != Phase::Accepted) continue;
            out.requests.push_back({e.from, e.to, to->bm_addr});
        }
        send(*c, out);
    }
}

void Master::handle_bandwidth_report(Client &c, const C2MReportBandwidth &p) {
    const Client *to = find(p.to_peer);
    if (!to || to->group != c.group) {
        PLOG(WARN) << "ignoring bandwidth report for unknown / foreign peer";
        return;
    }
    groups_[c.group].bandwidth.store(c.uuid, p.to_peer, p.mbits_per_second);
}

void Master::handle_optimize_work_complete(Client &c) {
    if (c.phase != Phase::Accepted || c.state != State::OptimizeTopology)
        return kick(c.link, "optimize-work-complete vote in illegal state");
    c.state = State::VoteCompleteTopologyOptimization;
    eval_optimize_complete();
}

void Master::run_ring_optimisation(uint32_t group) {
    Group &g = groups_[group];
    if (g.moonshot.valid()) g.moonshot.wait();
    {
        std::lock_guard<std::mutex> lock(topo_mu_);
        if (g.ring_optimal) return;
    }
    if (!g.optimised_once) {
        // first pass: solved before the peers are released
        g.optimised_once = true;
        std::vector<Uuid> ring;
        {
            std::lock_guard<std::mutex> lock(topo_mu_);
            ring = g.ring;
        }
        bool optimal = false;
        if (topo::optimize_ring(g.bandwidth, ring, optimal)) {
            std::lock_guard<std::mutex> lock(topo_mu_);
            g.ring = ring;
            g.ring_optimal = optimal;
        } else {
            PLOG(WARN) << "ring optimisation found no feasible ring for group " << group << "\n"
                       << g.bandwidth.debug_matrix();
        }
        return;
    }
    // later passes: long-running improvement in the background, picked up by the next round
    topo::BandwidthStore snapshot = g.bandwidth;
    std::vector<Uuid> ring;
    {
        std::lock_guard<std::mutex> lock(topo_mu_);
        ring = g.ring;
    }
    Group *gp = &g; // map nodes are stable; the group is not erased before this task finished
    g.moonshot = topo_pool_.submit([this, gp, snapshot, ring]() mutable {
        bool optimal = false, improved = false;
        if (!topo::improve_ring_moonshot(snapshot, ring, optimal, improved)) return;
        std::lock_guard<std::mutex> lock(topo_mu_);
        Group &gg = *gp;
        if (gg.ring.size() != ring.size()) return; // membership changed meanwhile
        if (improved) gg.ring = ring;
        gg.ring_optimal = optimal;
    });
}

void Master::eval_optimize_complete() {
    size_t votes = 0;
    for (const auto &[u, c] : clients_) {
        if (c.phase == Phase::Registered) continue;
        if (c.state != State::VoteCompleteTopologyOptimization) return;
        ++votes;
    }
    if (votes == 0) return;
    // Edges nobody could measure are remembered as unreachable so later rounds do not stall on them.
    for (const auto &[u, c] : clients_) {
        if (c.phase != Phase::Accepted) continue;
        for (const auto &e : missing_edges(c)) unreachable_[e.from].insert(e.to);
    }
    std::set<uint32_t> touched;
    for (auto &[u, c] : clients_) {
        if (c.phase != Phase::Accepted) continue;
        c.state = State::Idle;
        touched.insert(c.group);
    }
    for (uint32_t group : touched) {
        ring_for(group, false); // refresh membership first
        run_ring_optimisation(group);
    }
    for (Client *c : ordered_clients()) {
        if (c->phase != Phase::Accepted) continue;
        M2COptimizeTopologyComplete out;
        out.success = true;
        out.ring_order = ring_for(c->group, false);
        send(*c, out);
    }
}

// ------------------------------------------------------------------------------------------
// shared state
// ------------------------------------------------------------------------------------------

void Master::handle_sync_shared_state(Client &c, const C2MSyncSharedState &p) {
    if (c.phase != Phase::Accepted || c.state != State::Idle) return kick(c.link, "shared-state vote in illegal state");
    Group &g = groups_[c.group];
    RevisionStatus rs = RevisionStatus::Match;
    if (g.revision_initialised) {
        if (p.revision < g.next_revision)
            rs = RevisionStatus::Outdated;
        else if (p.revision > g.next_revision)
            return kick(c.link, "shared-state revision must grow by exactly one per synchronisation");
    }
    // (The very first synchronisation of a group may start at any revision -- resume from a checkpoint.  Which one
    // that is gets decided when all votes are in, see eval_sync_votes.)
    if (static_cast<uint8_t>(p.strategy) > 2) return kick(c.link, "unknown sync strategy");
    c.state = State::VoteSyncSharedState;
    g.sync_votes.push_back({c.uuid, p.strategy, rs, p.entries, p.revision});
    if (eval_sync_votes(c.group) == Verdict::Failed) {
        PLOG(WARN) << "shared-state synchronisation of group " << c.group << " cannot proceed, kicking the group";
        for (Client *peer : ordered_clients())
            if (peer->group == c.group && peer->phase == Phase::Accepted) kick(peer->link, "shared-state sync failed");
    }
}

namespace {
// Peers of the reference build leave num_elements at zero (ccoip_client_handler.cpp:411-415 never fills it in), so a
// zero on either side means "not reported" rather than "empty tensor".
bool same_shape(const SharedStateEntry &a, const SharedStateEntry &b) {
    return a.key == b.key && a.allow_content_inequality == b.allow_content_inequality && a.dtype == b.dtype &&
           a.hash_type == b.hash_type && (a.num_elements == b.num_elements || a.num_elements == 0 || b.num_elements == 0);
}
bool same_candidate(const std::vector<SharedStateEntry> &a, const std::vector<SharedStateEntry> &b) {
    if (a.size() != b.size()) return false;
    for (size_t i = 0; i < a.size(); ++i)
        if (!same_shape(a[i], b[i]) || a[i].hash != b[i].hash) return false;
    return true;
}
} // namespace

Master::Verdict Master::eval_sync_votes(uint32_t group) {
    Group &g = groups_[group];
    size_t accepted = 0, voting = 0;
    for (const auto &[u, c] : clients_) {
        if (c.group != group || c.phase != Phase::Accepted) continue;
        ++accepted;
        if (c.state == State::VoteSyncSharedState) ++voting;
    }
    if (voting == 0 || voting != accepted || g.sync_votes.size() != voting) return Verdict::NotReady;

    // First synchronisation of the group: the highest revision anybody brings is where the group starts, and peers
    // below it are outdated.  (The reference fixes the starting revision to whatever the first non-zero VOTE carried,
    // ccoip_master_state.cpp:1076-1086, so a resumed peer and fresh peers could end up either way round depending on
    // packet arrival order; deciding with all votes on the table keeps the wire protocol and removes the race.)
    if (!g.revision_initialised) {
        uint64_t start = 0;
        for (const auto &v : g.sync_votes) start = std::max(start, v.revision);
        g.next_revision = start;
        for (auto &v : g.sync_votes) v.rev_status = v.revision < start ? RevisionStatus::Outdated : RevisionStatus::Match;
    }

    // ---- elect the reference content ("mask"): most votes among the candidates of the best
    // revision class; RECEIVE_ONLY peers do not stand for election.
    struct Tally {
        const std::vector<SharedStateEntry> *entries;
        size_t votes;
        int priority;
    };
    // Peers that mix another strategy into a group with an ENFORCE_POPULAR member are kicked further down; their content
    // does not stand for election either.  (The reference lets it: with one enforcer and one SEND_ONLY peer holding
    // different content the outcome -- enforcer served, or enforcer left without a distributor -- depends on whose vote
    // arrives first, ccoip_master_handler.cpp:616-735.)
    bool enforcer_present = false;
    for (const auto &v : g.sync_votes) enforcer_present = enforcer_present || v.strategy == SyncStrategy::EnforcePopular;
    std::vector<Tally> tallies;
    for (const auto &v : g.sync_votes) {
        if (v.strategy == SyncStrategy::RxOnly) continue;
        if (enforcer_present && v.strategy != SyncStrategy::EnforcePopular) continue;
        const int prio = v.rev_status == RevisionStatus::Match ? 1 : -2;
        auto it = std::find_if(tallies.begin(), tallies.end(), [&](const Tally &t) { return same_candidate(*t.entries, v.entries); });
        if (it == tallies.end())
            tallies.push_back({&v.entries, 1, prio});
        else {
            it->votes++;
            it->priority = std::max(it->priority, prio);
        }
    }
    // a tie in class and votes is broken by content (smaller hash sequence first), never by the order the votes came in
    auto content_less = [](const std::vector<SharedStateEntry> &a, const std::vector<SharedStateEntry> &b) {
        if (a.size() != b.size()) return a.size() < b.size();
        for (size_t i = 0; i < a.size(); ++i)
            if (a[i].hash != b[i].hash) return a[i].hash < b[i].hash;
        return false;
    };
    const Tally *winner = nullptr;
    for (const auto &t : tallies) {
        if (!winner || t.priority > winner->priority || (t.priority == winner->priority && t.votes > winner->votes) ||
            (t.priority == winner->priority && t.votes == winner->votes && content_less(*t.entries, *winner->entries)))
            winner = &t;
    }
    if (!winner) return Verdict::Failed; // nobody offered content
    const std::vector<SharedStateEntry> &mask = *winner->entries;

    // ---- classify every voter against the mask
    struct Outcome {
        ContentStatus status = ContentStatus::Match;
        std::vector<size_t> dirty; // indices into mask
    };
    std::unordered_map<Uuid, Outcome> outcomes;
    bool any_kick = false;
    for (const auto &v : g.sync_votes) {
        Outcome o;
        if (v.entries.size() != mask.size()) {
            o.status = ContentStatus::KeySetMismatch;
        } else {
            for (size_t i = 0; i < mask.size(); ++i) {
                if (!same_shape(mask[i], v.entries[i])) {
                    o.status = ContentStatus::KeySetMismatch;
                    break;
                }
                if (mask[i].hash != v.entries[i].hash) {
                    o.status = ContentStatus::ContentMismatch;
                    o.dirty.push_back(i);
                }
            }
        }
        if (o.status == ContentStatus::KeySetMismatch) {
            if (const Client *c = find(v.uuid)) {
                kick(c->link, "shared-state keys / types differ from the elected state");
                any_kick = true;
            }
        }
        outcomes.emplace(v.uuid, std::move(o));
    }
    // the kicked peers' disconnects re-trigger this evaluation with their votes removed
    if (any_kick) return Verdict::Done;

    // ---- assign roles
    std::unordered_set<Uuid> kicked;
    bool any_enforce = false, any_other = false;
    for (const auto &v : g.sync_votes) {
        Client *c = find(v.uuid);
        if (!c) continue;
        const Outcome &o = outcomes[v.uuid];
        const bool up_to_date = o.status == ContentStatus::Match && v.rev_status == RevisionStatus::Match;
        c->state = up_to_date ? State::DistributeSharedState : State::RequestSharedState;
        if (c->state == State::RequestSharedState && v.strategy == SyncStrategy::TxOnly) {
            kick(c->link, "declared SEND_ONLY but holds unpopular content");
            kicked.insert(c->uuid);
        }
        (v.strategy == SyncStrategy::EnforcePopular ? any_enforce : any_other) = true;
    }
    if (any_enforce && any_other) {
        for (const auto &v : g.sync_votes) {
            if (v.strategy == SyncStrategy::EnforcePopular) continue;
            if (Client *c = find(v.uuid)) {
                kick(c->link, "ENFORCE_POPULAR cannot be combined with other strategies in one group");
                kicked.insert(c->uuid);
            }
        }
    }
    g.next_revision++;
    g.revision_initialised = true;

    // ---- tell everybody.  Whoever needs content is pointed at a peer that has it: one on the same machine when there is
    // one (NVLink or loopback instead of the network), and among equally good ones the one with the fewest takers so far, so
    // that several newcomers do not all queue at the same peer.  (The reference always names the first distributor;
    // ccoip_master_handler.cpp:567-593 carries a "should be topology aware" note.)
    struct Source {
        const Client *peer;
        size_t takers = 0;
    };
    std::vector<Source> sources;
    for (Client *c : ordered_clients())
        if (c->group == group && c->phase == Phase::Accepted && c->state == State::DistributeSharedState && !kicked.count(c->uuid))
            sources.push_back({c, 0});
    auto same_machine = [](const Client &a, const Client &b) {
        const auto &x = a.ss_addr.inet, &y = b.ss_addr.inet;
        if (x.protocol != y.protocol) return false;
        return x.protocol == inetIPv4 ? std::memcmp(x.ipv4.data, y.ipv4.data, 4) == 0 : std::memcmp(x.ipv6.data, y.ipv6.data, 16) == 0;
    };
    auto pick_source = [&](const Client &taker) -> const Client * {
        Source *best = nullptr;
        bool best_local = false;
        for (auto &src : sources) {
            const bool local = same_machine(*src.peer, taker);
            if (!best || (local && !best_local) || (local == best_local && src.takers < best->takers)) best = &src, best_local = local;
        }
        if (!best) return nullptr;
        best->takers++;
        return best->peer;
    };
    for (Client *c : ordered_clients()) {
        if (c->group != group || c->phase != Phase::Accepted) continue;
        if (c->state != State::DistributeSharedState && c->state != State::RequestSharedState) continue;
        M2CSyncSharedState out;
        out.is_outdated = c->state == State::RequestSharedState;
        if (out.is_outdated) {
            const Client *distributor = kicked.count(c->uuid) ? nullptr : pick_source(*c);
            if (distributor) {
                out.distributor = distributor->ss_addr;
                for (size_t i : outcomes[c->uuid].dirty) {
                    out.outdated_keys.push_back(mask[i].key);
                    out.expected_hashes.push_back(mask[i].hash);
                    out.expected_hash_types.push_back(mask[i].hash_type);
                }
            } else if (!kicked.count(c->uuid)) {
                PLOG(WARN) << "no peer can distribute shared state to " << c->uuid.str();
            }
        }
        send(*c, out);
    }
    g.sync_votes.clear();
    return Verdict::Done;
}

void Master::handle_dist_complete(Client &c) {
    if (c.phase != Phase::Accepted ||
        (c.state != State::DistributeSharedState && c.state != State::RequestSharedState))
        return kick(c.link, "dist-complete vote in illegal state");
    c.state = State::VoteCompleteSharedStateSync;
    groups_[c.group].sync_complete_votes.insert(c.uuid);
    eval_sync_complete(c.group);
}

void Master::eval_sync_complete(uint32_t group) {
    Group &g = groups_[group];
    size_t accepted = 0, voting = 0;
    for (const auto &[u, c] : clients_) {
        if (c.group != group || c.phase != Phase::Accepted) continue;
        ++accepted;
        if (c.state == State::VoteCompleteSharedStateSync) ++voting;
    }
    if (voting == 0 || voting != accepted) return;
    g.sync_complete_votes.clear();
    for (Client *c : ordered_clients()) {
        if (c->group != group || c->phase != Phase::Accepted) continue;
        c->state = State::Idle;
        send(*c, M2CSyncSharedStateComplete{});
    }
}

// ------------------------------------------------------------------------------------------
// collectives
// ------------------------------------------------------------------------------------------

void Master::handle_collective_initiate(Client &c, const C2MCollectiveInitiate &p) {
    if (c.phase != Phase::Accepted || (c.state != State::Idle && c.state != State::CollectiveCommunicationsRunning))
        return kick(c.link, "collective-initiate vote in illegal state");
    if (const auto it = c.collectives.find(p.tag); it != c.collectives.end()) {
        // extension: a peer that already voted Complete may vote for the NEXT operation on the
        // same tag right away; the vote takes effect when the current one is finished
        if (it->second != CollectiveState::VoteComplete || c.queued_initiate.count(p.tag))
            return kick(c.link, "collective tag already in flight");
        c.queued_initiate[p.tag] = static_cast<uint8_t>(p.op) & 0xf8u;
        return;
    }
    // A vote for a tag the group runs under a lease: this peer cannot (or will not) take the lease path for this
    // operation.  The lease of that tag ends for everybody; whoever already started the operation without a vote
    // gives up and votes as well.
    if (groups_[c.group].leased_tags.count(p.tag)) revoke_lease(c.group, p.tag, false, false);
    c.state = State::CollectiveCommunicationsRunning;
    c.collectives[p.tag] = CollectiveState::VoteInitiate;
    // extension: top bit of the operator byte = "I can run this one on the device plane"
    // bit 6: "... and over NVSwitch multicast", bit 5: "I would run later operations on this tag under a lease",
    // bit 4: "the multicast view of the ring exists here" (independent of this operation)
    c.device_capable[p.tag] = static_cast<uint8_t>(p.op) & 0xf8u;
    eval_collective_initiate(c.group, p.tag);
}

void Master::eval_collective_initiate(uint32_t group, uint64_t tag) {
    size_t accepted = 0, voting = 0;
    for (const auto &[u, c] : clients_) {
        if (c.group != group || c.phase != Phase::Accepted) continue;
        ++accepted;
        const auto it = c.collectives.find(tag);
        if (it != c.collectives.end() && it->second == CollectiveState::VoteInitiate) ++voting;
    }
    if (voting == 0 || voting != accepted) return;
    uint64_t seq = next_seq_nr_++;
    // the device plane is used only if every single peer of the group offered it
    uint8_t common = 0xf8u;
    for (auto &[u, c] : clients_) {
        if (c.group != group || c.phase != Phase::Accepted) continue;
        const auto it = c.device_capable.find(tag);
        common &= it != c.device_capable.end() ? it->second : 0;
        c.device_capable.erase(tag);
    }
    if (common & 0x80u) seq |= 1ull << 63;
    if ((common & 0xc0u) == 0xc0u) seq |= 1ull << 62;
    if (common & 0x08u) seq |= 1ull << 59; // every peer has the two-level layout of a multi-host ring wired
    // Lease: every peer ran this operation on the device plane and offers to run the following operations on this
    // tag without voting.  From the next operation on, the group orders its collectives on this tag among itself
    // (same-box sequencer) and the master only hears of them when something goes wrong.
    if ((common & 0xa0u) == 0xa0u && accepted >= 2) {
        seq |= 1ull << 61;
        if (common & 0x10u) seq |= 1ull << 60;
        groups_[group].leased_tags.insert(tag);
    }
    for (Client *c : ordered_clients()) {
        if (c->group != group || c->phase != Phase::Accepted) continue;
        c->collectives[tag] = CollectiveState::Perform;
        M2CCollectiveCommence out;
        out.tag = tag;
        out.seq_nr = seq;
        send(*c, out);
    }
}

void Master::revoke_lease(uint32_t group, uint64_t tag, bool all, bool peer_lost) {
    Group &g = groups_[group];
    if (all ? g.leased_tags.empty() : !g.leased_tags.count(tag)) return;
    if (all)
        g.leased_tags.clear();
    else
        g.leased_tags.erase(tag);
    M2CLeaseRevoked out;
    out.tag = tag;
    out.all = all;
    out.peer_lost = peer_lost;
    for (Client *c : ordered_clients())
        if (c->group == group && c->phase == Phase::Accepted) send(*c, out);
}

bool Master::mark_aborted(uint32_t group, uint64_t tag) {
    bool &flag = groups_[group].aborted[tag];
    if (flag) return false;
    flag = true;
    return true;
}

void Master::broadcast_abort(uint32_t group, uint64_t tag, bool aborted) {
    M2CCollectiveAbort out;
    out.tag = tag;
    out.aborted = aborted;
    for (Client *c : ordered_clients()) {
        if (c->group != group || c->phase != Phase::Accepted) continue;
        if (!c->collectives.count(tag)) continue;
        send(*c, out);
    }
}

void Master::handle_collective_complete(Client &c, const C2MCollectiveComplete &p) {
    const auto it = c.collectives.find(p.tag);
    if (c.phase != Phase::Accepted || it == c.collectives.end() || it->second != CollectiveState::Perform)
        return kick(c.link, "collective-complete vote in illegal state");
    it->second = CollectiveState::VoteComplete;
    if (p.was_aborted && mark_aborted(c.group, p.tag)) broadcast_abort(c.group, p.tag, true);
    eval_collective_complete(c.group, p.tag);
}

void Master::eval_collective_complete(uint32_t group, uint64_t tag) {
    size_t accepted = 0, voting = 0;
    for (const auto &[u, c] : clients_) {
        if (c.group != group || c.phase != Phase::Accepted) continue;
        ++accepted;
        const auto it = c.collectives.find(tag);
        if (it != c.collectives.end() && it->second == CollectiveState::VoteComplete) ++voting;
    }
    if (voting == 0 || voting != accepted) return;
    Group &g = groups_[group];
    // Every operation yields exactly one Abort packet per peer: either the real abort while it
    // ran, or this "not aborted" notice at the end.  Peers can therefore always consume one.
    if (!g.aborted[tag]) broadcast_abort(group, tag, false);
    g.aborted.erase(tag);
    bool requeued = false;
    for (Client *c : ordered_clients()) {
        if (c->group != group || c->phase != Phase::Accepted) continue;
        c->collectives.erase(tag);
        if (c->collectives.empty()) c->state = State::Idle;
        M2CCollectiveComplete out;
        out.tag = tag;
        send(*c, out);
        if (const auto q = c->queued_initiate.find(tag); q != c->queued_initiate.end()) {
            c->state = State::CollectiveCommunicationsRunning;
            c->collectives[tag] = CollectiveState::VoteInitiate;
            c->device_capable[tag] = q->second;
            c->queued_initiate.erase(q);
            requeued = true;
        }
    }
    if (requeued) eval_collective_initiate(group, tag);
}

// ------------------------------------------------------------------------------------------
// disconnects
// ------------------------------------------------------------------------------------------

void Master::on_disconnect(const ccoip_socket_address_t &from) {
    state_owner_.check("master state machine");
    const auto ait = by_addr_.find(net::AddrKey(from));
    if (ait == by_addr_.end()) return; // never registered
    const Uuid uuid = ait->second;
    by_addr_.erase(ait);
    const auto cit = clients_.find(uuid);
    if (cit == clients_.end()) return;
    const uint32_t group = cit->second.group;
    const bool was_accepted = cit->second.phase == Phase::Accepted;
    std::set<uint64_t> tags;
    for (const auto &[u, c] : clients_)
        if (c.group == group)
            for (const auto &[tag, st] : c.collectives) tags.insert(tag);
    clients_.erase(cit);
    PLOG(INFO) << "peer " << uuid.str() << " left (" << net::to_string(from) << ")";

    Group &g = groups_[group];
    if (was_accepted) g.bandwidth.unregister_peer(uuid);
    if (was_accepted) revoke_lease(group, 0, true, true); // peers may be inside a leased operation waiting for the one that left
    g.sync_votes.erase(std::remove_if(g.sync_votes.begin(), g.sync_votes.end(),
                                      [&](const SyncVote &v) { return v.uuid == uuid; }),
                       g.sync_votes.end());
    g.sync_complete_votes.erase(uuid);
    last_p2p_info_.erase(uuid);
    unreachable_.erase(uuid);
    for (auto &[u, set] : unreachable_) set.erase(uuid);

    // Every pending decision may have become unanimous now that this voter is gone.
    eval_establish_votes();
    peer_dropped_ = true; // an establishment round in progress must be repeated
    eval_p2p_established();
    eval_peers_pending();
    eval_optimize_votes();
    eval_optimize_complete();
    if (eval_sync_votes(group) == Verdict::Failed) {
        for (Client *peer : ordered_clients())
            if (peer->group == group && peer->phase == Phase::Accepted) kick(peer->link, "shared-state sync failed");
    }
    eval_sync_complete(group);
    for (uint64_t tag : tags) {
        eval_collective_initiate(group, tag);
        // If everybody who is left had already finished, the operation completes normally.
        eval_collective_complete(group, tag);
        // Otherwise it cannot finish without the lost peer: tell the survivors to stop.
        bool running = false;
        for (const auto &[u, c] : clients_) {
            if (c.group != group) continue;
            const auto it = c.collectives.find(tag);
            if (it != c.collectives.end() && it->second == CollectiveState::Perform) running = true;
        }
        if (running && mark_aborted(group, tag)) broadcast_abort(group, tag, true);
    }
    if (local_world(group, true) == 0) {
        Group &gg = groups_[group];
        if (gg.moonshot.valid()) gg.moonshot.wait();
        groups_.erase(group);
    }
}

} // namespace pccl::master
