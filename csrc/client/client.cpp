#include "client.hpp"

#include <algorithm>
#include <chrono>
#include <cstring>
#include <future>
#include <random>
#include <thread>
#include <unistd.h>

#include "../kernels/cpu/hash.hpp"
#include "../kernels/cpu/reduce.hpp"
#include "../kernels/mset_formula.hpp"
#include "../net/sockutil.hpp"
#include "../util/env.hpp"
#include "../util/log.hpp"
#include "ring_reduce.hpp"

namespace pccl::client {

using namespace pccl::wire;
using clock_type = std::chrono::steady_clock;

namespace {
size_t frame_bytes_from_env() {
    // reference knob: PCCL_MULTIPLEX_CHUNK_SIZE (ccoip/src/cpp/reduce.cpp:20-41); smaller default
    // here because received frames are reduced while the next one is in flight
    return static_cast<size_t>(env_u64("PCCL_MULTIPLEX_CHUNK_SIZE", 2u << 20));
}
} // namespace

Client::Client(const ClientConfig &cfg)
    : cfg_(cfg), master_(cfg.master), p2p_listener_(net::any_like(cfg.master, cfg.p2p_port), true),
      ss_listener_(net::any_like(cfg.master, cfg.ss_port), true), bm_listener_(net::any_like(cfg.master, cfg.bm_port), true),
      plane_(device::Plane::create()) {
    if (cfg_.pool_size == 0) cfg_.pool_size = 1;
    const size_t n_workers = static_cast<size_t>(env_u64("PCCL_MAX_CONCURRENT_COLLECTIVE_OPS", 16));
    workers_ = std::make_unique<ThreadPool>(std::max<size_t>(1, n_workers), 64);
    ss_workers_ = std::make_unique<ThreadPool>(4, 64); // a requester may fetch over several connections at once
    bm_workers_ = std::make_unique<ThreadPool>(static_cast<size_t>(env_u64("PCCL_NUM_BENCHMARK_CONNECTIONS", 16)) + 1, 64);
    lease_enabled_ = env_flag("PCCL_LEASE", true);
    // which host this peer runs on, for the two-level all-reduce of groups that span hosts
    auto fnv = [](const std::string &text) {
        uint64_t h = 1469598103934665603ull;
        for (unsigned char c : text) h = (h ^ c) * 1099511628211ull;
        return h ? h : 1;
    };
    const char *host_env = std::getenv("PCCL_HOST_ID");
    if (cfg_.host_id) {
        host_id_ = cfg_.host_id;
    } else if (host_env && *host_env) {
        host_id_ = fnv(host_env);
    } else {
        char name[256] = {0};
        ::gethostname(name, sizeof name - 1);
        std::string id = name;
        if (FILE *f = std::fopen("/proc/sys/kernel/random/boot_id", "r")) {
            char boot[64] = {0};
            if (std::fgets(boot, sizeof boot, f)) id += boot;
            std::fclose(f);
        }
        host_id_ = fnv(id);
    }
    if (cfg_.hierarchy >= 0)
        hierarchy_enabled_ = cfg_.hierarchy != 0;
    else if (const char *v = std::getenv("PCCL_HIERARCHY"))
        hierarchy_enabled_ = v[0] != '0';
    else
        hierarchy_enabled_ = !net::is_loopback(cfg_.master) || cfg_.host_id != 0 || (host_env && *host_env);
}

Client::~Client() {
    interrupt();
    join();
}

// ------------------------------------------------------------------------------------------
// connect / lifecycle
// ------------------------------------------------------------------------------------------

bool Client::connect() {
    if (started_.exchange(true)) return false;
    p2p_listener_.on_join([this](std::unique_ptr<net::BlockingSocket> s) { on_p2p_accept(std::move(s)); });
    ss_listener_.on_join([this](std::unique_ptr<net::BlockingSocket> s) {
        auto shared = std::shared_ptr<net::BlockingSocket>(std::move(s));
        ss_workers_->submit([this, shared]() mutable {
            on_shared_state_accept(std::make_unique<net::BlockingSocket>(shared->release(), shared->remote()));
        });
    });
    bm_listener_.on_join([this](std::unique_ptr<net::BlockingSocket> s) {
        auto shared = std::shared_ptr<net::BlockingSocket>(std::move(s));
        bm_workers_->submit([this, shared]() mutable {
            on_benchmark_accept(std::make_unique<net::BlockingSocket>(shared->release(), shared->remote()));
        });
    });
    if (!p2p_listener_.listen() || !ss_listener_.listen() || !bm_listener_.listen()) return false;
    if (!p2p_listener_.run_async() || !ss_listener_.run_async() || !bm_listener_.run_async()) return false;

    if (!master_.connect()) {
        PLOG(ERR) << "cannot reach master at " << net::to_string(cfg_.master);
        return false;
    }
    // device-plane operations are queued per staging slot in the order their Commence packets arrive
    master_.set_observer([this](uint16_t id, const std::vector<uint8_t> &payload) {
        Reader r(payload.data(), payload.size());
        if (id == M2CLeaseRevoked::id) {
            // the lease of a tag (or of the whole group) ended: operations that run without a vote must stop waiting
            // for peers that will not come, and later operations vote again
            M2CLeaseRevoked rv;
            if (!rv.read(r)) return;
            {
                std::lock_guard<std::mutex> lock(lease_mu_);
                if (rv.all)
                    lease_tags_.clear();
                else
                    lease_tags_.erase(rv.tag);
            }
            if (rv.peer_lost) ring_stale_ = true;
            plane_->revoke_leased(rv.tag, rv.all);
            return;
        }
        if (id != M2CCollectiveCommence::id) return;
        M2CCollectiveCommence c;
        if (c.read(r) && (c.seq_nr >> 63)) plane_->announce(c.seq_nr & ~(31ull << 59));
    });
    master_.run();

    C2MSessionRegistration reg;
    reg.peer_group = cfg_.peer_group;
    reg.explicit_addresses = cfg_.explicit_addresses;
    if (cfg_.explicit_addresses) {
        reg.adv_p2p = cfg_.adv_p2p;
        reg.adv_ss = cfg_.adv_ss;
        reg.adv_bm = cfg_.adv_bm;
    } else {
        reg.p2p_port = p2p_listener_.port();
        reg.ss_port = ss_listener_.port();
        reg.bm_port = bm_listener_.port();
    }
    if (!master_.send_packet(reg)) return false;
    const auto resp = master_.recv_packet<M2CSessionRegistrationResponse>();
    if (!resp || !resp->accepted) {
        PLOG(ERR) << "master refused the registration";
        return false;
    }
    uuid_ = resp->assigned_uuid;
    plane_->start(uuid_);

    // Blocks until the accepted peers vote to let us in.
    const Establish first = establish_round();
    if (first == Establish::Failed) return false;
    accepted_ = true;
    if (first == Establish::Retry) {
        // We count as accepted even though the round has to be repeated (somebody dropped or was
        // unreachable); take part in the repeat like everybody else.
        if (!request_and_establish(true)) {
            accepted_ = false;
            return false;
        }
    } else {
        connection_revision_++;
    }
    return true;
}

bool Client::interrupt() {
    if (interrupted_.exchange(true)) return true;
    master_.interrupt();
    p2p_listener_.interrupt();
    ss_listener_.interrupt();
    bm_listener_.interrupt();
    {
        std::lock_guard<std::mutex> lock(tx_mu_);
        for (auto &[u, pool] : tx_pools_)
            for (auto &s : pool)
                if (s) s->interrupt();
    }
    {
        std::lock_guard<std::mutex> lock(rx_mu_);
        for (auto &[u, pool] : rx_pools_)
            for (auto &s : pool)
                if (s) s->interrupt();
    }
    rx_cv_.notify_all();
    return true;
}

bool Client::join() {
    if (workers_) workers_->shutdown();
    master_.join();
    p2p_listener_.join();
    ss_listener_.join();
    bm_listener_.join();
    if (ss_workers_) ss_workers_->shutdown();
    if (bm_workers_) bm_workers_->shutdown();
    {
        std::lock_guard<std::mutex> lock(tx_mu_);
        tx_pools_.clear();
    }
    {
        std::lock_guard<std::mutex> lock(rx_mu_);
        rx_pools_.clear();
    }
    {
        std::lock_guard<std::mutex> lock(stage_mu_);
        for (auto &[tag, b] : stage_cache_) plane_->host_free(b.ptr);
        stage_cache_.clear();
    }
    plane_->stop();
    return true;
}

// ------------------------------------------------------------------------------------------
// P2P wiring
// ------------------------------------------------------------------------------------------

void Client::on_p2p_accept(std::unique_ptr<net::BlockingSocket> sock) {
    const auto hello = sock->recv_packet<P2PHello>(5000);
    if (!hello) return;
    if (!sock->send_packet(P2PHelloAck{})) return;
    const ccoip_socket_address_t remote = sock->remote();
    auto mux = std::make_shared<net::MuxSocket>(sock->release(), remote, net::MuxSocket::Mode::Rx);
    mux->run();
    {
        std::lock_guard<std::mutex> lock(rx_mu_);
        Pool &pool = rx_pools_[hello->uuid];
        if (pool.size() <= hello->connection_nr) pool.resize(hello->connection_nr + 1);
        if (pool[hello->connection_nr]) pool[hello->connection_nr]->interrupt(); // replaced by a fresh link
        pool[hello->connection_nr] = std::move(mux);
    }
    rx_cv_.notify_all();
}

bool Client::dial_peer(const PeerEndpoint &peer) {
    Pool pool;
    for (uint32_t nr = 0; nr < cfg_.pool_size; ++nr) {
        net::BlockingSocket sock(peer.p2p_addr);
        if (!sock.connect(5000)) return false;
        P2PHello hello;
        hello.uuid = uuid_;
        hello.connection_nr = nr;
        if (!sock.send_packet(hello)) return false;
        if (!sock.recv_packet<P2PHelloAck>(5000)) return false;
        auto mux = std::make_shared<net::MuxSocket>(sock.release(), peer.p2p_addr, net::MuxSocket::Mode::Tx);
        mux->run();
        pool.push_back(std::move(mux));
    }
    std::lock_guard<std::mutex> lock(tx_mu_);
    tx_pools_[peer.uuid] = std::move(pool);
    return true;
}

Client::Establish Client::establish_round() {
    const auto info = master_.recv_packet<M2CP2PConnectionInfo>();
    if (!info) return Establish::Failed;
    PLOG(DEBUG) << "connection info: " << info->peers.size() << " neighbours, unchanged=" << info->unchanged;
    global_world_ = info->global_world_size;
    local_world_ = info->local_world_size;
    num_groups_ = info->num_peer_groups;
    largest_group_ = info->largest_group_size;

    C2MP2PEstablished report;
    report.success = true;
    if (!info->unchanged) last_neighbours_ = info->peers;
    if (!info->unchanged) {
        for (const auto &peer : info->peers) {
            bool have = false;
            {
                std::lock_guard<std::mutex> lock(tx_mu_);
                const auto it = tx_pools_.find(peer.uuid);
                have = it != tx_pools_.end() && !it->second.empty() &&
                       std::all_of(it->second.begin(), it->second.end(), [](const auto &s) { return s && s->is_open(); });
            }
            if (have) continue;
            if (!dial_peer(peer)) {
                PLOG(WARN) << "cannot reach peer " << peer.uuid.str() << " at " << net::to_string(peer.p2p_addr);
                report.success = false;
                report.failed_peers.push_back(peer.uuid);
                break;
            }
        }
        // links to peers that are no longer neighbours are closed
        auto wanted = [&](const Uuid &u) {
            return std::any_of(info->peers.begin(), info->peers.end(), [&](const PeerEndpoint &p) { return p.uuid == u; });
        };
        {
            std::lock_guard<std::mutex> lock(tx_mu_);
            for (auto it = tx_pools_.begin(); it != tx_pools_.end();) {
                if (wanted(it->first)) {
                    ++it;
                    continue;
                }
                for (auto &s : it->second)
                    if (s) s->interrupt();
                it = tx_pools_.erase(it);
            }
        }
        {
            std::lock_guard<std::mutex> lock(rx_mu_);
            for (auto it = rx_pools_.begin(); it != rx_pools_.end();) {
                if (wanted(it->first)) {
                    ++it;
                    continue;
                }
                for (auto &s : it->second)
                    if (s) s->interrupt();
                it = rx_pools_.erase(it);
            }
        }
    }
    PLOG(DEBUG) << "p2p links " << (report.success ? "established" : "FAILED") << ", reporting to the master";
    if (!master_.send_packet(report)) return Establish::Failed;
    const auto verdict = master_.recv_packet<M2CP2PEstablished>();
    PLOG(DEBUG) << "master verdict on p2p establishment received";
    if (!verdict) return Establish::Failed;
    if (!report.success && verdict->success) return Establish::Failed;
    if (!verdict->success) return Establish::Retry;
    {
        std::unique_lock<std::shared_mutex> lock(ring_mu_);
        ring_ = verdict->ring_order;
    }
    local_world_ = verdict->ring_order.size();
    ring_stale_ = false;
    {
        std::lock_guard<std::mutex> lock(lease_mu_); // leases belong to a ring
        lease_tags_.clear();
        lease_counts_.clear();
    }
    plane_->on_ring_changed(verdict->ring_order);
    build_hierarchy(verdict->ring_order);
    return Establish::Success;
}

// Two-level layout for rings that span hosts.  The peers first learn each other's host and address: a P2PDirectory packet
// travels around the ring like an all-gather (world-1 hops), every peer contributing its own host identity and the endpoint
// of its successor (the master only tells a peer about its neighbours).  If the ring turns out to be B >= 2 hosts with the
// same number L >= 2 of peers each, links to the local successor and to the peer of equal local rank on the next host are
// dialled.  Whether an operation then uses the hierarchy is agreed through the vote (bit 3 of the operator byte), so a peer
// on which any of this failed -- or a reference peer, which does not forward the directory -- simply keeps the ring flat
// for everybody.  The flat ring remains wired.  Quantised operations use the hierarchy too: plain values inside a host, codes
// between hosts.
void Client::build_hierarchy(const std::vector<Uuid> &ring) {
    {
        std::unique_lock<std::shared_mutex> lock(ring_mu_);
        hier_ = Hierarchy{};
    }
    const uint32_t W = static_cast<uint32_t>(ring.size());
    if (!hierarchy_enabled_ || W < 4) return;
    const auto me = std::find(ring.begin(), ring.end(), uuid_);
    if (me == ring.end()) return;
    const uint32_t r = static_cast<uint32_t>(me - ring.begin());
    const Uuid &next = ring[(r + 1) % W], &prev = ring[(r + W - 1) % W];
    Pool tx;
    {
        std::lock_guard<std::mutex> lock(tx_mu_);
        const auto it = tx_pools_.find(next);
        if (it != tx_pools_.end()) tx = it->second;
    }
    const Pool rx = pick_rx_pool(prev, 3000);
    if (tx.empty() || rx.empty()) return;

    constexpr uint64_t kDirectoryTag = 0xD17EC7027A6ull;
    uint64_t seq = 1469598103934665603ull; // the same on every member of this ring, different for another ring
    for (const auto &u : ring)
        for (uint8_t b : u.data) seq = (seq ^ b) * 1099511628211ull;
    seq &= (1ull << 59) - 1;

    std::unordered_map<Uuid, DirectoryEntry> known;
    auto merge = [&](const DirectoryEntry &e) {
        DirectoryEntry &k = known[e.uuid];
        k.uuid = e.uuid;
        if (e.has_host) {
            k.has_host = true;
            k.host_id = e.host_id;
        }
        if (e.has_endpoint) {
            k.has_endpoint = true;
            k.endpoint = e.endpoint;
        }
    };
    P2PDirectory mine;
    {
        DirectoryEntry self;
        self.uuid = uuid_;
        self.has_host = true;
        self.host_id = host_id_;
        mine.entries.push_back(self);
        for (const auto &p : last_neighbours_)
            if (p.uuid == next) {
                DirectoryEntry succ;
                succ.uuid = next;
                succ.has_endpoint = true;
                succ.endpoint = p.p2p_addr;
                mine.entries.push_back(succ);
            }
    }
    for (const auto &e : mine.entries) merge(e);
    P2PDirectory cur = mine;
    // the packet makes W-1 hops, each of them a one-way trip between two peers that may be an ocean apart
    const auto deadline = clock_type::now() + std::chrono::seconds(3) + std::chrono::milliseconds(150) * W;
    for (uint32_t step = 0; step + 1 < W; ++step) {
        if (!tx[0]->send_packet(kDirectoryTag, seq, cur)) return;
        std::optional<P2PDirectory> got;
        while (!got) {
            got = rx[0]->recv_packet<P2PDirectory>(kDirectoryTag, seq, 20);
            if (!got && (clock_type::now() > deadline || interrupted_ || !rx[0]->is_open())) {
                PLOG(INFO) << "peer directory did not come round (a peer of another build in the ring?): the ring stays flat";
                return;
            }
        }
        for (const auto &e : got->entries) merge(e);
        cur = std::move(*got);
    }

    // hosts in the order in which the ring first visits them; every host must hold the same number of peers
    std::vector<uint64_t> hosts;
    std::unordered_map<uint64_t, std::vector<Uuid>> members;
    for (const auto &u : ring) {
        const auto it = known.find(u);
        if (it == known.end() || !it->second.has_host) return;
        if (!members.count(it->second.host_id)) hosts.push_back(it->second.host_id);
        members[it->second.host_id].push_back(u);
    }
    const uint32_t B = static_cast<uint32_t>(hosts.size());
    const uint32_t L = static_cast<uint32_t>(members[hosts[0]].size());
    if (B < 2 || L < 2) return;
    for (uint64_t h : hosts)
        if (members[h].size() != L) {
            PLOG(INFO) << "hosts hold different numbers of peers: the ring stays flat";
            return;
        }
    Hierarchy h;
    h.B = B;
    h.L = L;
    h.box = static_cast<uint32_t>(std::find(hosts.begin(), hosts.end(), host_id_) - hosts.begin());
    if (h.box >= B) return;
    const auto &local = members[host_id_];
    h.local_rank = static_cast<uint32_t>(std::find(local.begin(), local.end(), uuid_) - local.begin());
    h.local_next = local[(h.local_rank + 1) % L];
    h.local_prev = local[(h.local_rank + L - 1) % L];
    h.cross_next = members[hosts[(h.box + 1) % B]][h.local_rank];
    h.cross_prev = members[hosts[(h.box + B - 1) % B]][h.local_rank];
    for (const Uuid *target : {&h.local_next, &h.cross_next}) {
        // A link to a ring neighbour is shared with the flat ring and was just confirmed by the establishment round.  Any
        // other link left over from an earlier layout is dialled afresh: the far end may have dropped its half in a round
        // in which this peer's neighbours did not change (so nothing was pruned here), and a sending socket does not notice
        // that nobody is listening until it writes.
        const bool ring_neighbour = std::any_of(last_neighbours_.begin(), last_neighbours_.end(), [&](const PeerEndpoint &p) { return p.uuid == *target; });
        bool have = false;
        if (ring_neighbour) {
            std::lock_guard<std::mutex> lock(tx_mu_);
            const auto it = tx_pools_.find(*target);
            have = it != tx_pools_.end() && !it->second.empty() &&
                   std::all_of(it->second.begin(), it->second.end(), [](const auto &s) { return s && s->is_open(); });
        }
        if (have) continue;
        const auto it = known.find(*target);
        if (it == known.end() || !it->second.has_endpoint) return;
        PeerEndpoint ep;
        ep.uuid = *target;
        ep.p2p_addr = it->second.endpoint;
        if (!dial_peer(ep)) {
            PLOG(WARN) << "cannot reach " << target->str() << " for the two-level ring: the ring stays flat";
            return;
        }
    }
    h.ready = true;
    PLOG(INFO) << "two-level ring: " << B << " hosts x " << L << " peers, this peer is (" << h.box << ", " << h.local_rank << ")";
    std::unique_lock<std::shared_mutex> lock(ring_mu_);
    hier_ = h;
}

bool Client::request_and_establish(bool accept_new_peers) {
    if (!accepted_ || !master_.is_open()) return false;
    if (any_reduce_running()) return false;
    drain_deferred(); // the master must see this peer idle before it may vote
    Establish result;
    do {
        C2MEstablishP2P vote;
        vote.accept_new_peers = accept_new_peers;
        if (!master_.send_packet(vote)) return false;
        result = establish_round();
    } while (result == Establish::Retry);
    if (result == Establish::Failed) return false;
    connection_revision_++;
    return true;
}

bool Client::update_topology() {
    std::lock_guard<std::mutex> lock(joint_mu_);
    return request_and_establish(true);
}

bool Client::are_peers_pending(bool &pending) {
    if (!accepted_ || !master_.is_open()) return false;
    if (!master_.send_packet(C2MCheckPeersPending{})) return false;
    const auto resp = master_.recv_packet<M2CPeersPendingResponse>();
    if (!resp) return false;
    pending = resp->pending;
    return true;
}

// ------------------------------------------------------------------------------------------
// topology optimisation
// ------------------------------------------------------------------------------------------

void Client::on_benchmark_accept(std::unique_ptr<net::BlockingSocket> sock) {
    const auto hello = sock->recv_packet<C2BHello>(5000);
    if (!hello) return;
    // One measuring peer at a time; the runner opens several connections, so further
    // connections of the admitted peer are let in, everybody else is told to come back.
    bool admitted = false;
    {
        std::lock_guard<std::mutex> lock(bm_mu_);
        if (bm_active_ == 0 || bm_owner_ == hello->uuid) {
            bm_owner_ = hello->uuid;
            bm_active_++;
            admitted = true;
        }
    }
    B2CBusy reply;
    reply.busy = !admitted;
    sock->send_packet(reply);
    if (!admitted) return;
    std::vector<uint8_t> sink(1 << 20);
    const auto deadline = clock_type::now() + std::chrono::seconds(18);
    while (clock_type::now() < deadline && !interrupted_) {
        const int rc = net::wait_readable(sock->fd(), 200);
        if (rc < 0) break;
        if (rc == 0) continue;
        const ssize_t k = ::recv(sock->fd(), sink.data(), sink.size(), 0);
        if (k <= 0) break;
    }
    std::lock_guard<std::mutex> lock(bm_mu_);
    --bm_active_;
}

bool Client::run_benchmark(const BenchmarkRequest &req, double &mbits, bool &busy) {
    busy = false;
    const int n_conn = static_cast<int>(env_u64("PCCL_NUM_BENCHMARK_CONNECTIONS", 16));
    const uint64_t millis = env_u64("PCCL_BENCHMARK_MILLIS", 10000);
    std::vector<std::unique_ptr<net::BlockingSocket>> socks;
    for (int i = 0; i < n_conn; ++i) {
        auto s = std::make_unique<net::BlockingSocket>(req.endpoint);
        if (!s->connect(3000)) return false;
        C2BHello hello;
        hello.uuid = uuid_;
        if (!s->send_packet(hello)) return false;
        const auto reply = s->recv_packet<B2CBusy>(5000);
        if (!reply) return false;
        if (reply->busy) {
            busy = true;
            return false;
        }
        socks.push_back(std::move(s));
    }
    std::vector<uint8_t> payload(8u << 20);
    std::mt19937_64 rng(0xbeef);
    for (size_t i = 0; i + 8 <= payload.size(); i += 8) {
        const uint64_t v = rng();
        std::memcpy(&payload[i], &v, 8);
    }
    std::atomic<uint64_t> total{0};
    const auto start = clock_type::now();
    const auto stop_at = start + std::chrono::milliseconds(millis);
    std::vector<std::thread> threads;
    for (auto &s : socks) {
        threads.emplace_back([&, fd = s->fd()] {
            net::set_send_timeout(fd, 2);
            while (clock_type::now() < stop_at && !interrupted_) {
                const ssize_t k = ::send(fd, payload.data(), std::min<size_t>(payload.size(), 1u << 20), MSG_NOSIGNAL);
                if (k <= 0) break;
                total += static_cast<uint64_t>(k);
            }
        });
    }
    for (auto &t : threads) t.join();
    const double secs = std::chrono::duration<double>(clock_type::now() - start).count();
    socks.clear();
    mbits = secs > 0 ? static_cast<double>(total) * 8.0 / 1e6 / secs : 0.0;
    return total > 0;
}

bool Client::optimize_topology() {
    std::lock_guard<std::mutex> lock(joint_mu_);
    if (!accepted_ || !master_.is_open()) return false;
    if (any_reduce_running()) return false;
    drain_deferred();
    for (int round = 0; round < 64; ++round) {
        if (!master_.send_packet(C2MOptimizeTopology{})) return false;
        const auto work = master_.recv_packet<M2COptimizeTopologyResponse>();
        if (!work) return false;
        std::vector<BenchmarkRequest> todo = work->requests;
        int attempts = 0;
        while (!todo.empty() && attempts < 30 && !interrupted_) {
            std::vector<BenchmarkRequest> later;
            for (const auto &req : todo) {
                double mbits = 0;
                bool busy = false;
                if (run_benchmark(req, mbits, busy)) {
                    C2MReportBandwidth rep;
                    rep.to_peer = req.to;
                    rep.mbits_per_second = mbits;
                    if (!master_.send_packet(rep)) return false;
                    PLOG(DEBUG) << "bandwidth to " << req.to.str() << ": " << mbits << " Mbit/s";
                } else if (busy) {
                    later.push_back(req);
                    std::this_thread::sleep_for(std::chrono::milliseconds(env_u64("PCCL_BENCHMARK_BACKOFF_MILLIS", 1000)));
                } // unreachable endpoints are simply not reported; the master marks them
            }
            todo.swap(later);
            ++attempts;
        }
        if (!master_.send_packet(C2MOptimizeWorkComplete{})) return false;
        const auto done = master_.recv_packet<M2COptimizeTopologyComplete>();
        if (!done) return false;
        if (done->success) break;
    }
    // rewire to the neighbours of the new ring
    return request_and_establish(false);
}

// ------------------------------------------------------------------------------------------
// shared state
// ------------------------------------------------------------------------------------------

namespace {
constexpr size_t kStageBytes = 8u << 20;

// Whether the peer of a connected socket is this machine (the kernel picked the same address for both ends).
bool same_host(int fd) {
    sockaddr_storage a{}, b{};
    socklen_t la = sizeof a, lb = sizeof b;
    if (getsockname(fd, reinterpret_cast<sockaddr *>(&a), &la) != 0 || getpeername(fd, reinterpret_cast<sockaddr *>(&b), &lb) != 0) return true;
    if (a.ss_family != b.ss_family) return false;
    if (a.ss_family == AF_INET)
        return reinterpret_cast<sockaddr_in *>(&a)->sin_addr.s_addr == reinterpret_cast<sockaddr_in *>(&b)->sin_addr.s_addr;
    return std::memcmp(&reinterpret_cast<sockaddr_in6 *>(&a)->sin6_addr, &reinterpret_cast<sockaddr_in6 *>(&b)->sin6_addr, 16) == 0;
}

// two page-locked staging buffers for the socket <-> GPU pipelines of the shared-state transfer
class PinnedPair {
public:
    explicit PinnedPair(device::Plane &plane) : plane_(plane) {}
    ~PinnedPair() {
        for (void *p : bufs_)
            if (p) plane_.host_free(p);
    }
    bool ok() {
        for (void *&p : bufs_)
            if (!p && !(p = plane_.host_alloc(kStageBytes))) return false;
        return true;
    }
    uint8_t *buf(size_t i) { return static_cast<uint8_t *>(bufs_[i & 1]); }

private:
    device::Plane &plane_;
    void *bufs_[2] = {nullptr, nullptr};
};
} // namespace

void Client::on_shared_state_accept(std::unique_ptr<net::BlockingSocket> sock) {
    uint16_t id = 0;
    std::vector<uint8_t> payload;
    if (!sock->recv_frame(id, payload, 10000)) return;
    Reader r(payload.data(), payload.size());
    std::vector<std::string> keys;
    Uuid requester;
    const bool extended = id == C2SRequestSharedStateP2P::id;
    if (extended) {
        C2SRequestSharedStateP2P req;
        if (!req.read(r)) return;
        keys = std::move(req.keys);
        requester = req.requester;
    } else if (id == C2SRequestSharedState::id) {
        C2SRequestSharedState req;
        if (!req.read(r)) return;
        keys = std::move(req.keys);
    } else {
        return;
    }
    S2CSharedStateResponseP2P resp;
    std::vector<SharedEntry> serve;
    {
        std::lock_guard<std::mutex> lock(ss_mu_);
        if (!ss_serving_) {
            resp.status = SharedStateStatus::NotInDistributionMode;
        } else {
            resp.status = SharedStateStatus::Success;
            resp.revision = ss_snapshot_.revision;
            for (const auto &key : keys) {
                const auto it = std::find_if(ss_snapshot_.entries.begin(), ss_snapshot_.entries.end(),
                                             [&](const SharedEntry &e) { return e.key == key; });
                if (it == ss_snapshot_.entries.end()) {
                    resp.status = SharedStateStatus::NotDistributed;
                    resp.entries.clear();
                    serve.clear();
                    break;
                }
                resp.entries.push_back({key, it->bytes()});
                serve.push_back(*it);
            }
        }
    }
    // NVLink path: the requester lives in the same NVLink domain and has this peer's arena mapped
    const size_t chunk = extended ? plane_->exchange_bytes() : 0;
    resp.p2p = extended && chunk > 0 && plane_->peer_mapped(requester) && env_flag("PCCL_SHARED_STATE_P2P", true);
    resp.distributor = uuid_;
    resp.chunk_bytes = chunk;
    if (extended) {
        if (!sock->send_packet(resp)) return;
    } else {
        S2CSharedStateResponse base;
        base.status = resp.status;
        base.revision = resp.revision;
        base.entries = resp.entries;
        if (!sock->send_packet(base)) return;
    }
    if (resp.status != SharedStateStatus::Success) return;
    // The buffers stay valid: the owner is blocked inside sync_shared_state until every peer has
    // reported completion.
    if (resp.p2p) {
        auto staging = plane_->exchange_lock(); // one requester at a time uses the staging area
        // Chunk number c of the whole transfer goes through staging slot c % window; a slot is staged again only when the
        // requester has acknowledged the chunk that used it last.  With two slots this peer stages chunk c+1 while the
        // requester still copies chunk c out over NVLink.
        const size_t window = std::max<size_t>(1, std::min<size_t>(plane_->exchange_slots(), 2));
        uint64_t sent = 0, acked = 0;
        auto take_ack = [&]() -> bool {
            if (!sock->recv_packet<C2SChunkAck>(60000)) return false;
            ++acked;
            return true;
        };
        for (const auto &e : serve) {
            const size_t bytes = e.bytes();
            for (size_t off = 0; off < bytes || (bytes == 0 && off == 0); off += chunk) {
                const size_t n = std::min(chunk, bytes - off);
                while (sent - acked >= window)
                    if (!take_ack()) return;
                if (n && !plane_->export_chunk(sent % window, static_cast<const uint8_t *>(e.data) + off, e.on_device, n)) return;
                if (!sock->send_packet(S2CChunkReady{})) return;
                ++sent;
                if (bytes == 0) break;
            }
            ss_tx_bytes_ += bytes;
        }
        while (acked < sent)
            if (!take_ack()) return;
        return;
    }
    // Base protocol: raw tensor bytes follow, unframed, in request order.  Device tensors go through two page-locked buffers:
    // while one is on its way into the socket, the next chunk is copied off the GPU into the other.
    PinnedPair stage(*plane_);
    for (const auto &e : serve) {
        const size_t bytes = e.bytes();
        if (!e.on_device) {
            if (!sock->send_raw(e.data, bytes)) return;
        } else {
            if (!stage.ok()) return;
            const auto *src = static_cast<const uint8_t *>(e.data);
            size_t filled = std::min(kStageBytes, bytes);
            if (filled && !plane_->copy_to_host(stage.buf(0), src, filled)) return;
            for (size_t off = 0, k = 0; off < bytes; ++k) {
                const size_t n = filled, next_off = off + n, next_n = std::min(kStageBytes, bytes - next_off);
                std::future<bool> next;
                if (next_n) next = std::async(std::launch::async, [&, k, next_off, next_n] { return plane_->copy_to_host(stage.buf((k + 1) & 1), src + next_off, next_n); });
                const bool sent = sock->send_raw(stage.buf(k & 1), n);
                const bool copied = !next.valid() || next.get();
                if (!sent || !copied) return;
                off = next_off;
                filled = next_n;
            }
        }
        ss_tx_bytes_ += bytes;
    }
}

bool Client::sync_shared_state(SharedStateArgs &state, uint64_t &tx_bytes, uint64_t &rx_bytes) {
    std::lock_guard<std::mutex> lock(joint_mu_);
    tx_bytes = rx_bytes = 0;
    if (!accepted_ || !master_.is_open()) return false;
    if (any_reduce_running()) return false;
    drain_deferred();

    // simplehash, like the reference (ccoip_client_handler.cpp:391-394); PCCL_SHARED_STATE_HASH=crc32 switches
    // the whole run to CRC32C (all peers must agree, the master compares the hash type too)
    static const HashType hash_type = [] {
        const char *v = std::getenv("PCCL_SHARED_STATE_HASH");
        if (v && (std::string(v) == "crc32" || std::string(v) == "crc32c")) return HashType::Crc32;
        if (v && (std::string(v) == "mset" || std::string(v) == "mset64")) return HashType::Mset64;
        return HashType::Simple;
    }();
    auto hash_of = [&](const SharedEntry &e, uint64_t &out) {
        if (hash_type == HashType::Mset64 && trust_hashes_ && recall_hash(e.data, e.bytes(), out)) {
            ++hash_skips_; // the kernel that last wrote this tensor hashed it on the way
            return true;
        }
        if (e.on_device) return plane_->hash_device(e.data, e.bytes(), hash_type, out);
        out = hash_type == HashType::Simple  ? cpu::simplehash(e.data, e.bytes())
              : hash_type == HashType::Crc32 ? cpu::crc32c(e.data, e.bytes())
                                             : cpu::mset64(e.data, e.bytes());
        return true;
    };
    C2MSyncSharedState vote;
    vote.revision = state.revision;
    vote.strategy = state.strategy;
    // device tensors whose hash is not already known are hashed in one batch (one wait for all of them)
    std::vector<device::Plane::HashJob> jobs;
    std::vector<size_t> job_entry;
    std::vector<uint64_t> hashes(state.entries.size(), 0);
    std::vector<bool> have(state.entries.size(), false);
    for (size_t i = 0; i < state.entries.size(); ++i) {
        const auto &e = state.entries[i];
        if (e.allow_content_inequality) continue;
        if (hash_type == HashType::Mset64 && trust_hashes_ && recall_hash(e.data, e.bytes(), hashes[i])) {
            ++hash_skips_;
            have[i] = true;
        } else if (e.on_device && state.entries.size() > 1) {
            jobs.push_back({e.data, e.bytes(), 0});
            job_entry.push_back(i);
        }
    }
    { // host tensors that still need a hash: several at a time
        std::vector<const void *> ptrs;
        std::vector<size_t> sizes, idx;
        for (size_t i = 0; i < state.entries.size(); ++i) {
            const auto &e = state.entries[i];
            if (e.allow_content_inequality || e.on_device || have[i]) continue;
            ptrs.push_back(e.data), sizes.push_back(e.bytes()), idx.push_back(i);
        }
        if (ptrs.size() > 1) {
            std::vector<uint64_t> out(ptrs.size(), 0);
            cpu::hash_many(ptrs.data(), sizes.data(), ptrs.size(), static_cast<int>(hash_type), out.data());
            for (size_t k = 0; k < idx.size(); ++k) hashes[idx[k]] = out[k], have[idx[k]] = true;
        }
    }
    if (!jobs.empty()) {
        if (plane_->hash_device_many(jobs, hash_type)) {
            for (size_t j = 0; j < jobs.size(); ++j) {
                hashes[job_entry[j]] = jobs[j].result;
                have[job_entry[j]] = true;
            }
        } // else: one by one below, which reports the tensor that fails
    }
    for (size_t i = 0; i < state.entries.size(); ++i) {
        const auto &e = state.entries[i];
        SharedStateEntry w;
        w.key = e.key;
        w.hash_type = hash_type;
        w.num_elements = e.count;
        w.dtype = e.dtype;
        w.allow_content_inequality = e.allow_content_inequality;
        if (have[i]) {
            w.hash = hashes[i];
        } else if (!e.allow_content_inequality && !hash_of(e, w.hash)) {
            PLOG(ERR) << "cannot hash tensor '" << e.key << "'";
            return false;
        }
        vote.entries.push_back(std::move(w));
    }

    // From here until completion the shared-state server may hand our tensors to other peers.
    {
        std::lock_guard<std::mutex> l(ss_mu_);
        ss_snapshot_ = state;
        ss_serving_ = true;
        ss_tx_bytes_ = 0;
    }
    struct Guard {
        Client *c;
        ~Guard() {
            std::lock_guard<std::mutex> l(c->ss_mu_);
            c->ss_serving_ = false;
        }
    } guard{this};

    if (!master_.send_packet(vote)) return false;
    const auto verdict = master_.recv_packet<M2CSyncSharedState>();
    if (!verdict) return false;

    if (verdict->is_outdated) {
        const auto &d = verdict->distributor;
        const bool empty_v4 = d.inet.protocol == inetIPv4 && d.port == 0;
        if (empty_v4) {
            PLOG(ERR) << "shared state is outdated but nobody can distribute it";
            return false;
        }
        net::BlockingSocket sock(verdict->distributor);
        if (!sock.connect(10000)) {
            PLOG(ERR) << "cannot reach shared-state distributor " << net::to_string(verdict->distributor);
            return false;
        }
        // One tensor of the plain stream: straight into host memory, or socket -> page-locked buffer -> GPU with the copy of
        // chunk k overlapping the receive of chunk k+1.
        auto receive_tensor = [&](net::BlockingSocket &from, PinnedPair &stage, const SharedEntry &e, size_t bytes) {
            if (!e.on_device) return from.recv_raw(e.data, bytes);
            if (!stage.ok()) return false;
            auto *dst = static_cast<uint8_t *>(e.data);
            std::future<bool> pending;
            size_t k = 0;
            for (size_t off = 0; off < bytes; off += kStageBytes, ++k) {
                const size_t n = std::min<size_t>(kStageBytes, bytes - off);
                const bool got = from.recv_raw(stage.buf(k & 1), n);
                if (pending.valid() && !pending.get()) return false; // the buffer received into next is free again
                if (!got) return false;
                pending = std::async(std::launch::async, [&, k, off, n] { return plane_->copy_to_device(dst + off, stage.buf(k & 1), n); });
            }
            return !pending.valid() || pending.get();
        };
        auto entry_of = [&](const std::string &key) -> const SharedEntry * {
            const auto it = std::find_if(state.entries.begin(), state.entries.end(), [&](const SharedEntry &e) { return e.key == key; });
            return it == state.entries.end() ? nullptr : &*it;
        };

        // Ask with the extended request when this peer has a device plane that could take the data
        // over NVLink; the distributor answers whether it can.  A distributor on another machine cannot.  A large transfer
        // that has to go through sockets is split over several connections: one TCP stream rarely fills a long fat pipe.  Every
        // connection asks for a contiguous run of the outdated tensors with the reference's plain request, so any distributor
        // can serve it.
        const bool remote = !same_host(sock.fd());
        const bool try_p2p = plane_->exchange_bytes() > 0 && !remote;
        uint64_t want_bytes = 0;
        for (const auto &key : verdict->outdated_keys)
            if (const SharedEntry *e = entry_of(key)) want_bytes += e->bytes();
        size_t n_streams = static_cast<size_t>(env_u64("PCCL_SHARED_STATE_CONNECTIONS", 4));
        n_streams = std::min<size_t>({std::max<size_t>(n_streams, 1), verdict->outdated_keys.size(), 16});
        if (want_bytes < (32u << 20)) n_streams = 1;
        if (!try_p2p && n_streams > 1) {
            // contiguous runs of about equal size
            std::vector<std::pair<size_t, size_t>> runs; // [first, last) into outdated_keys
            {
                const uint64_t per = (want_bytes + n_streams - 1) / n_streams;
                size_t first = 0;
                uint64_t acc = 0;
                for (size_t i = 0; i < verdict->outdated_keys.size(); ++i) {
                    const SharedEntry *e = entry_of(verdict->outdated_keys[i]);
                    if (!e) return false;
                    acc += e->bytes();
                    if (acc >= per && runs.size() + 1 < n_streams) {
                        runs.push_back({first, i + 1});
                        first = i + 1;
                        acc = 0;
                    }
                }
                if (first < verdict->outdated_keys.size()) runs.push_back({first, verdict->outdated_keys.size()});
            }
            PLOG(DEBUG) << "fetching " << want_bytes << " bytes of shared state over " << runs.size() << " connections";
            std::vector<uint64_t> revisions(runs.size(), 0), moved(runs.size(), 0);
            std::vector<char> good(runs.size(), 0);
            std::vector<std::thread> streams;
            for (size_t g = 0; g < runs.size(); ++g)
                streams.emplace_back([&, g] {
                    std::unique_ptr<net::BlockingSocket> own;
                    net::BlockingSocket *link = &sock;
                    if (g > 0) {
                        own = std::make_unique<net::BlockingSocket>(verdict->distributor);
                        if (!own->connect(10000)) return;
                        link = own.get();
                    }
                    C2SRequestSharedState req;
                    req.keys.assign(verdict->outdated_keys.begin() + runs[g].first, verdict->outdated_keys.begin() + runs[g].second);
                    if (!link->send_packet(req)) return;
                    const auto resp = link->recv_packet<S2CSharedStateResponse>(60000);
                    if (!resp || resp->status != SharedStateStatus::Success || resp->entries.size() != req.keys.size()) return;
                    revisions[g] = resp->revision;
                    PinnedPair stage(*plane_);
                    for (const auto &slice : resp->entries) {
                        const SharedEntry *e = entry_of(slice.key);
                        if (!e || e->bytes() != slice.size_bytes) return;
                        if (!receive_tensor(*link, stage, *e, slice.size_bytes)) return;
                        moved[g] += slice.size_bytes;
                    }
                    good[g] = 1;
                });
            for (auto &t : streams) t.join();
            for (size_t g = 0; g < runs.size(); ++g)
                if (!good[g] || revisions[g] != revisions[0]) {
                    PLOG(ERR) << "shared-state transfer over " << runs.size() << " connections failed on connection " << g;
                    return false;
                }
            state.revision = revisions[0];
            std::vector<std::pair<const SharedEntry *, uint64_t>> arrived;
            for (size_t i = 0; i < verdict->outdated_keys.size(); ++i) {
                const SharedEntry *e = entry_of(verdict->outdated_keys[i]);
                rx_bytes += e->bytes();
                forget_hashes(e->data, e->bytes());
                if (!e->allow_content_inequality && i < verdict->expected_hashes.size()) arrived.push_back({e, verdict->expected_hashes[i]});
            }
            if (!verify_received(arrived, hash_type)) return false;
            if (!master_.send_packet(C2MDistSharedStateComplete{})) return false;
            if (!master_.recv_packet<M2CSyncSharedStateComplete>()) return false;
            tx_bytes = ss_tx_bytes_;
            return true;
        }
        bool p2p = false;
        Uuid distributor;
        uint64_t chunk = 0;
        std::vector<SharedStateSlice> offered;
        if (try_p2p) {
            C2SRequestSharedStateP2P req;
            req.requester = uuid_;
            req.keys = verdict->outdated_keys;
            if (!sock.send_packet(req)) return false;
            const auto resp = sock.recv_packet<S2CSharedStateResponseP2P>(60000);
            if (!resp || resp->status != SharedStateStatus::Success) {
                PLOG(ERR) << "distributor refused the shared-state request";
                return false;
            }
            p2p = resp->p2p && resp->chunk_bytes > 0 && plane_->peer_mapped(resp->distributor);
            if (resp->p2p && !p2p) return false; // the distributor will wait for acks we cannot give
            distributor = resp->distributor;
            chunk = resp->chunk_bytes;
            offered = resp->entries;
            state.revision = resp->revision;
        } else {
            C2SRequestSharedState req;
            req.keys = verdict->outdated_keys;
            if (!sock.send_packet(req)) return false;
            const auto resp = sock.recv_packet<S2CSharedStateResponse>(60000);
            if (!resp || resp->status != SharedStateStatus::Success) {
                PLOG(ERR) << "distributor refused the shared-state request";
                return false;
            }
            offered = resp->entries;
            state.revision = resp->revision; // adopt the distributor's revision even if no bytes move
        }
        if (offered.size() != verdict->outdated_keys.size()) return false;
        PinnedPair stage(*plane_);
        const size_t window = std::max<size_t>(1, std::min<size_t>(plane_->exchange_slots(), 2));
        uint64_t chunk_no = 0; // counts the chunks of the whole transfer, like the distributor does
        unsigned imports_pending = 0, acks_owed = 0;
        std::vector<std::pair<const SharedEntry *, uint64_t>> to_verify;
        for (size_t i = 0; i < offered.size(); ++i) {
            const auto &slice = offered[i];
            const auto it = std::find_if(state.entries.begin(), state.entries.end(),
                                         [&](const SharedEntry &e) { return e.key == slice.key; });
            if (it == state.entries.end() || it->bytes() != slice.size_bytes) {
                PLOG(ERR) << "distributor offered '" << slice.key << "' with an unexpected size";
                return false;
            }
            if (p2p) {
                // The copy of a chunk is only enqueued when its Ready arrives; the acknowledgement (which lets the distributor
                // reuse the slot) follows once the copy has finished -- one chunk later, so that consecutive copies run back
                // to back on the stream while the next Ready is already on its way.
                for (size_t off = 0; off < slice.size_bytes || (slice.size_bytes == 0 && off == 0); off += chunk) {
                    const size_t n = std::min<size_t>(chunk, slice.size_bytes - off);
                    if (!sock.recv_packet<S2CChunkReady>(60000)) return false;
                    if (n) {
                        if (!plane_->import_chunk_begin(chunk_no % window, distributor, static_cast<uint8_t *>(it->data) + off, it->on_device, n)) return false;
                        ++imports_pending;
                    } else {
                        ++acks_owed; // an empty tensor: nothing to copy, acknowledged with the others
                    }
                    while (imports_pending > (window > 1 ? 1u : 0u)) {
                        if (!plane_->import_chunk_finish_oldest()) return false;
                        --imports_pending;
                        ++acks_owed;
                    }
                    for (; acks_owed; --acks_owed)
                        if (!sock.send_packet(C2SChunkAck{})) return false;
                    ++chunk_no;
                    if (slice.size_bytes == 0) break;
                }

            } else if (!receive_tensor(sock, stage, *it, slice.size_bytes)) {
                return false;
            }
            rx_bytes += slice.size_bytes;
            forget_hashes(it->data, it->bytes());
            if (!it->allow_content_inequality && i < verdict->expected_hashes.size()) to_verify.push_back({&*it, verdict->expected_hashes[i]});
        }
        // the last copies must have landed before anything is hashed
        while (imports_pending) {
            if (!plane_->import_chunk_finish_oldest()) return false;
            --imports_pending;
            ++acks_owed;
        }
        for (; acks_owed; --acks_owed)
            if (!sock.send_packet(C2SChunkAck{})) return false;
        if (!verify_received(to_verify, hash_type)) return false;
    }

    if (!master_.send_packet(C2MDistSharedStateComplete{})) return false;
    if (!master_.recv_packet<M2CSyncSharedStateComplete>()) return false;
    tx_bytes = ss_tx_bytes_;
    return true;
}

// What arrived is checked against what the master elected: device tensors in one batch of hash kernels.
bool Client::verify_received(const std::vector<std::pair<const SharedEntry *, uint64_t>> &to_verify, wire::HashType hash_type) {
    auto hash_of = [&](const SharedEntry &e, uint64_t &out) { // nothing to recall here: whatever was known about these tensors is stale
        if (e.on_device) return plane_->hash_device(e.data, e.bytes(), hash_type, out);
        out = hash_type == HashType::Simple  ? cpu::simplehash(e.data, e.bytes())
              : hash_type == HashType::Crc32 ? cpu::crc32c(e.data, e.bytes())
                                             : cpu::mset64(e.data, e.bytes());
        return true;
    };
    {
        std::vector<device::Plane::HashJob> vjobs;
        std::vector<size_t> vidx;
        for (size_t k = 0; k < to_verify.size(); ++k)
            if (to_verify[k].first->on_device && to_verify.size() > 1) {
                vjobs.push_back({to_verify[k].first->data, to_verify[k].first->bytes(), 0});
                vidx.push_back(k);
            }
        std::vector<uint64_t> got(to_verify.size(), 0);
        std::vector<bool> hashed(to_verify.size(), false);
        if (!vjobs.empty() && plane_->hash_device_many(vjobs, hash_type))
            for (size_t j = 0; j < vjobs.size(); ++j) {
                got[vidx[j]] = vjobs[j].result;
                hashed[vidx[j]] = true;
            }
        {
            std::vector<const void *> ptrs;
            std::vector<size_t> sizes, idx;
            for (size_t k = 0; k < to_verify.size(); ++k)
                if (!hashed[k] && !to_verify[k].first->on_device) ptrs.push_back(to_verify[k].first->data), sizes.push_back(to_verify[k].first->bytes()), idx.push_back(k);
            if (ptrs.size() > 1) {
                std::vector<uint64_t> out(ptrs.size(), 0);
                cpu::hash_many(ptrs.data(), sizes.data(), ptrs.size(), static_cast<int>(hash_type), out.data());
                for (size_t j = 0; j < idx.size(); ++j) got[idx[j]] = out[j], hashed[idx[j]] = true;
            }
        }
        for (size_t k = 0; k < to_verify.size(); ++k) {
            const SharedEntry &e = *to_verify[k].first;
            if (!hashed[k] && !hash_of(e, got[k])) return false;
            if (got[k] != to_verify[k].second) {
                PLOG(ERR) << "shared-state tensor '" << e.key << "' does not match the elected hash after transfer";
                return false;
            }
            if (hash_type == HashType::Mset64) remember_hash(e.data, e.bytes(), got[k]);
        }
    }
    return true;
}

// ------------------------------------------------------------------------------------------
// all-reduce
// ------------------------------------------------------------------------------------------

void Client::remember_hash(const void *ptr, size_t bytes, uint64_t value) {
    std::lock_guard<std::mutex> lock(hash_mu_);
    known_hashes_[reinterpret_cast<uintptr_t>(ptr)] = KnownHash{bytes, value};
    if (known_hashes_.size() > 4096) known_hashes_.erase(known_hashes_.begin());
}

void Client::forget_hashes(const void *ptr, size_t bytes) {
    const auto lo = reinterpret_cast<uintptr_t>(ptr), hi = lo + bytes;
    std::lock_guard<std::mutex> lock(hash_mu_);
    for (auto it = known_hashes_.begin(); it != known_hashes_.end();) {
        if (it->first < hi && it->first + it->second.bytes > lo)
            it = known_hashes_.erase(it);
        else
            ++it;
    }
}

bool Client::recall_hash(const void *ptr, size_t bytes, uint64_t &value) {
    std::lock_guard<std::mutex> lock(hash_mu_);
    const auto it = known_hashes_.find(reinterpret_cast<uintptr_t>(ptr));
    if (it == known_hashes_.end() || it->second.bytes != bytes) return false;
    value = it->second.value;
    return true;
}

bool Client::any_reduce_running() {
    std::lock_guard<std::mutex> lock(ops_mu_);
    return !ops_.empty() || !stream_ops_.empty();
}

bool Client::last_stats(uint64_t tag, pcclExtReduceStats_t &out) {
    std::lock_guard<std::mutex> lock(ops_mu_);
    const auto it = stats_.find(tag);
    if (it == stats_.end()) return false;
    out = it->second;
    return true;
}

Client::Launch Client::all_reduce_async(const ReduceRequest &req) {
    if (!accepted_ || !master_.is_open()) return Launch::NotConnected;
    std::unique_lock<std::mutex> lock(ops_mu_);
    if (ops_.count(req.tag)) return Launch::DuplicateTag;
    Op op;
    op.revision_at_launch = connection_revision_;
    const uint64_t rev = op.revision_at_launch;
    bool leased_device = false;
    if (lease_enabled_ && !req.outer) {
        {
            std::lock_guard<std::mutex> l(lease_mu_);
            leased_device = lease_tags_.count(req.tag) > 0;
        }
        leased_device = leased_device && device::classify_pointer(req.send).is_device && device::classify_pointer(req.recv).is_device;
    }
    if (leased_device) {
        // no vote, no socket: the launch takes microseconds, do it here (ordered after the legacy default stream, like every
        // operation that is not given a stream) and let await_reduce() take the verdict
        ops_.emplace(req.tag, Op{});
        ReduceRequest inline_req = req;
        inline_req.user_stream = reinterpret_cast<void *>(uintptr_t(1)); // cudaStreamLegacy
        inline_req.enqueue = true;
        lock.unlock();
        ReduceOutcome out = run_reduce(inline_req, rev);
        lock.lock();
        Op &slot = ops_[req.tag];
        slot.revision_at_launch = rev;
        if (out.pending) {
            slot.launched_inline = true;
            slot.ticket = out.ticket;
            slot.partial = out;
        } else { // the lease was gone after all and the operation ran (or failed) right here
            std::promise<ReduceOutcome> done;
            done.set_value(out);
            slot.future = done.get_future();
        }
        return Launch::Ok;
    }
    op.future = workers_->submit([this, req, rev] { return run_reduce(req, rev); });
    ops_.emplace(req.tag, std::move(op));
    return Launch::Ok;
}

bool Client::await_reduce(uint64_t tag, ReduceOutcome &out) {
    Op op;
    {
        std::lock_guard<std::mutex> lock(ops_mu_);
        const auto it = ops_.find(tag);
        if (it == ops_.end() || (!it->second.future.valid() && !it->second.launched_inline)) return false;
        op = std::move(it->second);
    }
    if (op.launched_inline) {
        StreamOp so;
        so.ticket = op.ticket;
        so.revision_at_launch = op.revision_at_launch;
        so.partial = op.partial;
        {
            std::lock_guard<std::mutex> lock(ops_mu_);
            ops_.erase(tag);
        }
        return finish_stream_op(tag, so, out);
    }
    try {
        out = op.future.get();
    } catch (...) {
        out = ReduceOutcome{};
    }
    return conclude(tag, op.revision_at_launch, out);
}

Client::Launch Client::all_reduce_sync(const ReduceRequest &req, ReduceOutcome &out) {
    if (!accepted_ || !master_.is_open()) return Launch::NotConnected;
    uint64_t rev;
    {
        std::lock_guard<std::mutex> lock(ops_mu_);
        if (ops_.count(req.tag)) return Launch::DuplicateTag;
        Op op;
        op.revision_at_launch = rev = connection_revision_;
        ops_.emplace(req.tag, std::move(op));
    }
    out = run_reduce(req, rev);
    conclude(req.tag, rev, out);
    return Launch::Ok;
}

Client::Launch Client::all_reduce_on_stream(const ReduceRequest &req, ReduceOutcome &out) {
    if (!accepted_ || !master_.is_open()) return Launch::NotConnected;
    // Verdicts of earlier enqueued operations are collected on the way: all of them when too many are in flight (each
    // holds a flag bank), otherwise only while the oldest ones have already finished.  A failure among them is
    // reported by this call.
    constexpr size_t kMaxEnqueued = 8;
    for (;;) {
        StreamOp oldest;
        uint64_t tag = 0;
        {
            std::lock_guard<std::mutex> lock(ops_mu_);
            if (stream_ops_.size() < kMaxEnqueued) break;
            tag = stream_ops_.front().first;
            oldest = stream_ops_.front().second;
            stream_ops_.pop_front();
        }
        ReduceOutcome prev;
        if (!finish_stream_op(tag, oldest, prev)) {
            out = prev;
            out.ok = false;
            return Launch::Ok;
        }
    }
    uint64_t rev;
    {
        std::lock_guard<std::mutex> lock(ops_mu_);
        if (ops_.count(req.tag)) return Launch::DuplicateTag;
        Op op;
        op.revision_at_launch = rev = connection_revision_;
        ops_.emplace(req.tag, std::move(op));
    }
    out = run_reduce(req, rev);
    if (out.pending) {
        std::lock_guard<std::mutex> lock(ops_mu_);
        ops_.erase(req.tag); // several enqueued operations may share a tag; they are tracked in stream_ops_
        StreamOp so;
        so.ticket = out.ticket;
        so.revision_at_launch = rev;
        so.partial = out;
        stream_ops_.emplace_back(req.tag, so);
        return Launch::Ok;
    }
    conclude(req.tag, rev, out);
    return Launch::Ok;
}

bool Client::finish_stream_op(uint64_t tag, const StreamOp &so, ReduceOutcome &out) {
    const auto r = plane_->finish(so.ticket, [&]() { return !master_.is_open() || interrupted_.load(); });
    out = so.partial;
    out.pending = false;
    out.ok = r.ok;
    out.tx_bytes = r.tx_bytes;
    out.rx_bytes = r.rx_bytes;
    out.stats.plane = r.plane;
    out.stats.device_ms = r.device_ms;
    out.stats.kernel_launches = r.kernel_launches;
        std::copy(std::begin(r.phase_ns), std::end(r.phase_ns), std::begin(out.stats.phase_ns));
    out.has_hash = r.has_hash;
    out.hash[0] = r.hash[0];
    out.hash[1] = r.hash[1];
    if (!r.ok) {
        std::lock_guard<std::mutex> lock(lease_mu_);
        lease_tags_.erase(tag);
    }
    bool others;
    {
        std::lock_guard<std::mutex> lock(ops_mu_);
        stats_[tag] = out.stats;
        others = !ops_.empty() || !stream_ops_.empty();
    }
    if (!out.ok && !others) {
        std::lock_guard<std::mutex> lock(joint_mu_);
        if (connection_revision_ == so.revision_at_launch && master_.is_open()) request_and_establish(false);
    }
    return out.ok;
}

bool Client::wait_stream_ops(uint64_t tag, bool all, ReduceOutcome &last) {
    std::vector<std::pair<uint64_t, StreamOp>> todo;
    {
        std::lock_guard<std::mutex> lock(ops_mu_);
        for (auto it = stream_ops_.begin(); it != stream_ops_.end();) {
            if (all || it->first == tag) {
                todo.push_back(*it);
                it = stream_ops_.erase(it);
            } else {
                ++it;
            }
        }
    }
    bool ok = true;
    for (auto &[t, so] : todo) ok = finish_stream_op(t, so, last) && ok;
    return ok;
}

bool Client::conclude(uint64_t tag, uint64_t revision_at_launch, ReduceOutcome &out) {
    bool others_running;
    {
        std::lock_guard<std::mutex> lock(ops_mu_);
        ops_.erase(tag);
        stats_[tag] = out.stats;
        if (stats_.size() > 4096) stats_.erase(stats_.begin());
        others_running = !ops_.empty();
    }
    if (!out.ok) {
        PLOG(DEBUG) << "reduction on tag " << tag << " failed, re-establishing p2p links";
        // A peer dropped: fetch the new ring before the caller retries.  Only once per
        // connection revision, and only when no other reduction is still in flight (those
        // will fail too and the last one to be awaited does the rewiring).
        std::lock_guard<std::mutex> lock(joint_mu_);
        if (!others_running && connection_revision_ == revision_at_launch && master_.is_open())
            request_and_establish(false);
    }
    return out.ok;
}

// Consumes what the master still owes for an operation whose Complete vote was already sent:
// exactly one Abort packet, then the Complete packet.
bool Client::finish_deferred(uint64_t tag, Deferred &d) {
    if (!d.abort_seen) {
        const auto pkt = master_.recv_matching<M2CCollectiveAbort>([&](const M2CCollectiveAbort &p) { return p.tag == tag; });
        if (!pkt) return false;
        d.abort_seen = true;
        d.aborted = pkt->aborted;
    }
    return master_.recv_matching<M2CCollectiveComplete>([&](const M2CCollectiveComplete &p) { return p.tag == tag; }).has_value();
}

bool Client::drain_deferred() {
    std::map<uint64_t, Deferred> pending;
    {
        std::lock_guard<std::mutex> lock(deferred_mu_);
        pending.swap(deferred_);
    }
    bool ok = true;
    for (auto &[tag, d] : pending) {
        ok = finish_deferred(tag, d) && ok;
        if (d.aborted) ring_stale_ = true;
    }
    return ok;
}

Client::Pool Client::pick_rx_pool(const Uuid &peer, int timeout_ms) {
    std::unique_lock<std::mutex> lock(rx_mu_);
    auto deadline = clock_type::now() + std::chrono::milliseconds(timeout_ms);
    bool saw_dead_link = false;
    for (;;) {
        const auto it = rx_pools_.find(peer);
        if (it != rx_pools_.end() && it->second.size() >= cfg_.pool_size) {
            bool all = true;
            for (uint32_t i = 0; i < cfg_.pool_size; ++i) all = all && it->second[i] && it->second[i]->is_open();
            if (all) return Pool(it->second.begin(), it->second.begin() + cfg_.pool_size);
            // Links that exist but are closed mean the predecessor died (or is re-dialling right
            // now): give a re-dial a moment, but do not sit out the full timeout meant for links
            // that have not been accepted yet.
            if (!saw_dead_link) {
                saw_dead_link = true;
                deadline = std::min(deadline, clock_type::now() + std::chrono::milliseconds(250));
            }
        }
        if (interrupted_ || rx_cv_.wait_until(lock, deadline) == std::cv_status::timeout) return {};
    }
}

uint8_t *Client::staging_for(uint64_t tag, size_t bytes) {
    std::lock_guard<std::mutex> lock(stage_mu_);
    StageBuf &b = stage_cache_[tag];
    if (b.bytes < bytes) {
        if (b.ptr) plane_->host_free(b.ptr);
        b.ptr = static_cast<uint8_t *>(plane_->host_alloc(bytes));
        b.bytes = b.ptr ? bytes : 0;
    }
    return b.ptr;
}

ReduceOutcome Client::run_reduce(const ReduceRequest &req, uint64_t revision_at_launch) {
    ReduceOutcome out = run_reduce_impl(req, revision_at_launch);
    if (out.ok && !out.pending && out.has_hash) {
        // the kernel hashed what it wrote: the next shared-state synchronisation can use it (see trust_content_hashes)
        if (req.outer) {
            const size_t bytes = req.count * sizeof(float);
            const uint64_t hp = mset::finalize(out.hash[0], bytes), hm = mset::finalize(out.hash[1], bytes);
            remember_hash(req.outer->params, bytes, hp);
            remember_hash(req.outer->momentum, bytes, hm);
            if (req.outer->inner) remember_hash(req.outer->inner, bytes, hp); // inner receives the new parameters
            out.stats.content_hash[0] = hp;
            out.stats.content_hash[1] = hm;
        } else {
            const size_t bytes = req.count * dtype_size(req.dtype);
            out.stats.content_hash[0] = mset::finalize(out.hash[0], bytes);
            remember_hash(req.recv, bytes, out.stats.content_hash[0]);
        }
        out.stats.has_content_hash = 1;
    }
    return out;
}

ReduceOutcome Client::run_reduce_impl(const ReduceRequest &req, uint64_t) {
    ReduceOutcome out;
    const size_t esz = dtype_size(req.dtype);
    // Buffers of the reduction proper.  For an outer step they are either unused (fused kernel) or a scratch
    // buffer holding the pseudo-gradient (two-pass form).
    const void *send = req.outer ? req.outer->params : req.send;
    void *recv = req.outer ? req.outer->params : req.recv;
    const auto send_info = device::classify_pointer(send), recv_info = device::classify_pointer(recv);

    // whatever this operation writes no longer has a known content hash
    if (req.outer) {
        forget_hashes(req.outer->params, req.count * sizeof(float));
        forget_hashes(req.outer->momentum, req.count * sizeof(float));
        if (req.outer->inner) forget_hashes(req.outer->inner, req.count * sizeof(float));
    } else {
        forget_hashes(req.recv, req.count * esz);
    }
    std::vector<Uuid> ring;
    {
        std::shared_lock<std::shared_mutex> lock(ring_mu_);
        ring = ring_;
    }
    const auto me = std::find(ring.begin(), ring.end(), uuid_);
    if (me == ring.end()) return out;
    const uint32_t world = static_cast<uint32_t>(ring.size()), rank = static_cast<uint32_t>(me - ring.begin());
    out.world = world;

    device::DeviceReduceArgs dargs;
    dargs.src = send;
    dargs.dst = recv;
    dargs.count = req.count;
    dargs.dtype = req.dtype;
    dargs.op = req.op;
    dargs.quant = req.quant;
    dargs.quant_dtype = req.quant_dtype;
    dargs.rank = rank;
    dargs.world = world;
    dargs.outer = req.outer;
    dargs.user_stream = req.user_stream;
    dargs.want_hash = req.want_hash || (trust_hashes_ && req.outer != nullptr);
    uint32_t caps = send_info.is_device && recv_info.is_device ? plane_->capabilities(dargs) : 0;
    // Outer step that cannot run as one kernel (host tensors, no device plane, larger than a staging slot): the
    // pseudo-gradient goes to a scratch buffer now, the scratch buffer is reduced like any tensor, and the
    // update is applied once the verdict is in.  Every peer sees the same geometry and decides alike.
    const bool outer_two_pass = req.outer && (caps & device::Plane::kCapP2P) == 0;
    std::vector<float> outer_host_scratch;
    std::unique_lock<std::mutex> scratch_lock;
    if (outer_two_pass) {
        dargs.outer = nullptr;
        dargs.want_hash = false; // the reduction's result is the averaged pseudo-gradient, not a tensor anybody synchronises
        float *scratch = nullptr;
        if (send_info.is_device) {
            scratch_lock = plane_->fuse_lock();
            scratch = static_cast<float *>(plane_->fuse_buffer(req.count * sizeof(float)));
            if (!scratch || !plane_->pseudo_grad(scratch, req.outer->params, req.outer->inner, req.count)) {
                PLOG(ERR) << "outer step: cannot stage the pseudo-gradient on the device";
                return out;
            }
        } else {
            outer_host_scratch.resize(req.count);
            for (uint64_t i = 0; i < req.count; ++i) outer_host_scratch[i] = req.outer->params[i] - req.outer->inner[i];
            scratch = outer_host_scratch.data();
        }
        send = recv = scratch;
        dargs.src = dargs.dst = scratch;
        caps = send_info.is_device ? plane_->capabilities(dargs) : 0;
    }
    const bool outer_fused = req.outer && !outer_two_pass;
    auto apply_outer_update = [&]() -> bool {
        if (!outer_two_pass) return true;
        const auto *avg = static_cast<const float *>(recv);
        if (send_info.is_device) return plane_->outer_nesterov(*req.outer, avg, req.count);
        const device::OuterStepArgs &o = *req.outer;
        for (uint64_t i = 0; i < req.count; ++i) { // torch.optim.SGD(nesterov=True), as in the device kernels
            const float g = avg[i];
            const float m = o.first ? g : o.mu * o.momentum[i] + g;
            const float p = o.params[i] - o.lr * (g + o.mu * m);
            o.momentum[i] = m;
            o.params[i] = p;
            if (o.inner) o.inner[i] = p;
        }
        return true;
    };
    const bool device_capable = (caps & device::Plane::kCapP2P) != 0;

    // ---- lease path.  An earlier operation on this tag was voted, ran on the device plane on every peer, and its
    // Commence carried the lease bit: from then on the group runs this tag's collectives without the master.  Their
    // order among concurrently issued operations comes from the same-box sequencer (device/sequencer.hpp), their
    // failure detection from the master all the same: it revokes the lease when a member vanishes or when somebody
    // votes on the tag after all (M2CLeaseRevoked, handled on the receive thread), which makes the kernels give up.
    if (lease_enabled_) {
        bool leased = false;
        uint64_t k = 0;
        {
            std::lock_guard<std::mutex> lock(lease_mu_);
            const auto it = lease_tags_.find(req.tag);
            if (it != lease_tags_.end()) {
                if (device_capable && !ring_stale_.load()) {
                    leased = true;
                    k = lease_counts_[req.tag]++;
                } else {
                    lease_tags_.erase(it); // host buffers, unsupported type...: vote; the master ends the lease for the others
                }
            }
        }
        if (leased) {
            dargs.leased = true;
            dargs.tag = req.tag;
            dargs.lease_k = k;
            dargs.use_nvls = lease_mc_all_.load() && (caps & device::Plane::kCapNvls) != 0;
            dargs.enqueue_only = req.enqueue && req.user_stream != nullptr && !outer_two_pass;
            const auto r = plane_->reduce(dargs, [&]() { return !master_.is_open() || interrupted_.load(); });
            if (r.pending) { // launched behind the caller's stream; the verdict is taken later (wait_stream_ops)
                out.pending = true;
                out.ticket = r.ticket;
                out.stats.leased = 1;
                out.stats.seq_nr = k;
                out.ok = true;
                return out;
            }
            out.has_hash = r.has_hash;
            out.hash[0] = r.hash[0];
            out.hash[1] = r.hash[1];
            out.tx_bytes = r.tx_bytes;
            out.rx_bytes = r.rx_bytes;
            out.stats.plane = r.plane;
            out.stats.device_ms = r.device_ms;
            out.stats.kernel_launches = r.kernel_launches;
        std::copy(std::begin(r.phase_ns), std::end(r.phase_ns), std::begin(out.stats.phase_ns));
            out.stats.seq_nr = k;
            out.stats.leased = 1;
            if (r.ok) {
                out.ok = apply_outer_update();
                return out;
            }
            {
                std::lock_guard<std::mutex> lock(lease_mu_);
                lease_tags_.erase(req.tag);
            }
            // The lease ended underneath the operation (not a data-plane failure) and the input is intact: vote like a
            // peer that never had the lease.  Anything else is reported; the caller rewires and retries.
            if (!(r.revoked && !r.input_destroyed && master_.is_open() && !ring_stale_.load())) return out;
            dargs.leased = false;
            out.stats = pcclExtReduceStats_t{};
        }
    }

    // ---- vote.  The spare top bit of the operator byte tells the master that this peer can
    // run the operation on the device plane; the master answers with the top bit of the
    // sequence number when EVERY peer of the group said so.  Masters and peers that do not
    // know the convention never set either bit, so mixed runs fall back to the TCP ring.
    C2MCollectiveInitiate vote;
    vote.tag = req.tag;
    vote.count = req.count;
    vote.dtype = req.dtype;
    // (bit 6 / bit 62 do the same for the NVSwitch multicast variant)
    // (bit 5: this peer would run LATER operations on this tag under a lease; bit 4: it holds the multicast view)
    const bool offer_lease = lease_enabled_ && device_capable && plane_->lease_ready();
    // (bit 3: this peer has the two-level layout of a multi-host ring wired and would use it for this operation)
    Hierarchy hier;
    {
        std::shared_lock<std::shared_mutex> lock(ring_mu_);
        hier = hier_;
    }
    const bool offer_hierarchy = hier.ready && hier.L * hier.B == world;
    vote.op = static_cast<RedOp>(static_cast<uint8_t>(req.op) | (device_capable ? 0x80u : 0u) |
                                 ((caps & device::Plane::kCapNvls) ? 0x40u : 0u) | (offer_lease ? 0x20u : 0u) |
                                 (offer_lease && plane_->multicast_ready() ? 0x10u : 0u) | (offer_hierarchy ? 0x08u : 0u));
    // The previous operation on this tag may have returned before its completion handshake was
    // consumed (device plane, see below).  A master of this build queues the new vote behind
    // it, so the vote goes out first and both answers arrive in one round trip; towards any
    // other master the handshake is finished before voting again.
    Deferred prev;
    bool have_prev = false;
    {
        std::lock_guard<std::mutex> lock(deferred_mu_);
        // Handshakes of OTHER tags whose answers have arrived in the meantime are retired on the
        // way (programs that use a fresh tag for every operation would otherwise pile them up).
        for (auto it = deferred_.begin(); it != deferred_.end();) {
            Deferred &d = it->second;
            const uint64_t tag = it->first;
            if (tag != req.tag) {
                if (!d.abort_seen) {
                    if (const auto pkt = master_.recv_matching<M2CCollectiveAbort>([&](const M2CCollectiveAbort &p) { return p.tag == tag; }, true)) {
                        d.abort_seen = true;
                        d.aborted = pkt->aborted;
                    }
                }
                if (d.abort_seen &&
                    master_.recv_matching<M2CCollectiveComplete>([&](const M2CCollectiveComplete &p) { return p.tag == tag; }, true)) {
                    if (d.aborted) ring_stale_ = true;
                    it = deferred_.erase(it);
                    continue;
                }
            }
            ++it;
        }
        const auto it = deferred_.find(req.tag);
        if (it != deferred_.end()) {
            prev = it->second;
            have_prev = true;
            deferred_.erase(it);
        }
    }
    if (have_prev && !master_pipelines_ && !finish_deferred(req.tag, prev)) return out;
    if (!master_.send_packet(vote)) return out;
    if (have_prev && master_pipelines_ && !finish_deferred(req.tag, prev)) return out;
    const auto go = master_.recv_matching<M2CCollectiveCommence>([&](const M2CCollectiveCommence &p) { return p.tag == req.tag; });
    if (!go) return out;
    // the earlier operation finished here, but a peer vanished before everybody had voted: the
    // ring this peer knows is stale, so this operation is reported as failed without running
    if (have_prev && prev.aborted) ring_stale_ = true;
    const bool poisoned = ring_stale_.load();
    const bool use_device = (go->seq_nr >> 63) != 0;
    const bool use_nvls = use_device && ((go->seq_nr >> 62) & 1) != 0;
    const uint64_t seq = go->seq_nr & ~(31ull << 59);
    out.stats.seq_nr = seq;
    const bool use_hierarchy = !use_device && offer_hierarchy && ((go->seq_nr >> 59) & 1) != 0; // every peer offered it
    // bit 61 also says that every peer holds the same-box sequencer: the operation is ordered through it
    dargs.sequenced = use_device && ((go->seq_nr >> 61) & 1) != 0;
    if (use_device && offer_lease && ((go->seq_nr >> 61) & 1) != 0) {
        // lease granted: the NEXT operation on this tag skips the vote (every peer decides so from this same packet)
        lease_mc_all_ = ((go->seq_nr >> 60) & 1) != 0;
        std::lock_guard<std::mutex> lock(lease_mu_);
        lease_tags_.insert(req.tag);
    }
    if (use_device) master_pipelines_ = true; // only a master of this build ever sets that bit

    bool abort_seen = false, aborted = false;
    auto abort_requested = [&]() {
        if (abort_seen) return aborted;
        const auto pkt = master_.recv_matching<M2CCollectiveAbort>([&](const M2CCollectiveAbort &p) { return p.tag == req.tag; }, true);
        if (pkt) {
            abort_seen = true;
            aborted = pkt->aborted;
        }
        if (!master_.is_open()) return true;
        return aborted;
    };

    bool local_ok = false;
    // in-place rollback for host buffers: the tag's cached vector when the cache has room (a tag is in flight once, so nobody
    // else touches it meanwhile), otherwise one that lives as long as this call
    std::vector<uint8_t> uncached_backup;
    std::vector<uint8_t> *cached = nullptr;
    if (send == recv && !send_info.is_device && !req.input_is_scratch) {
        constexpr uint64_t kCacheLimit = 4ull << 30;
        const uint64_t bytes = req.count * esz;
        std::lock_guard<std::mutex> lock(stage_mu_);
        auto it = backup_cache_.find(req.tag);
        if (it == backup_cache_.end() && backup_cache_bytes_ + bytes <= kCacheLimit) {
            it = backup_cache_.emplace(req.tag, std::vector<uint8_t>()).first;
            backup_cache_bytes_ += bytes;
        }
        if (it != backup_cache_.end() && (it->second.capacity() >= bytes || backup_cache_bytes_ + bytes <= kCacheLimit)) cached = &it->second;
    }
    std::vector<uint8_t> &host_backup = cached ? *cached : uncached_backup;
    host_backup.clear(); // a left-over of the previous operation on this tag must not pass for this one's backup
    uint8_t *stage = nullptr;         // host staging for device buffers on the TCP plane
    const bool in_place = send == recv && !outer_fused;

    if (use_device && (poisoned || !device_capable)) plane_->cancel(seq); // announced, but it will not run
    if (poisoned) {
        PLOG(INFO) << "previous reduction on tag " << req.tag << " was aborted after it finished locally; failing this one";
    } else if (use_device && device_capable) {
        dargs.seq = seq;
        dargs.use_nvls = use_nvls;
        const auto r = plane_->reduce(dargs, abort_requested);
        local_ok = r.ok;
        out.tx_bytes = r.tx_bytes;
        out.rx_bytes = r.rx_bytes;
        out.stats.plane = r.plane;
        out.stats.device_ms = r.device_ms;
        out.stats.kernel_launches = r.kernel_launches;
        std::copy(std::begin(r.phase_ns), std::end(r.phase_ns), std::begin(out.stats.phase_ns));
        out.has_hash = r.has_hash;
        out.hash[0] = r.hash[0];
        out.hash[1] = r.hash[1];
    } else if (use_device) {
        // cannot happen unless peers disagree about their own capability between vote and run
        PLOG(ERR) << "device plane was agreed on but is not available locally";
    } else {
        out.stats.plane = pcclExtPlaneTcpRing;
        RingContext ctx;
        ctx.tag = req.tag;
        ctx.seq = seq;
        ctx.rank = rank;
        ctx.world = world;
        ctx.frame_bytes = frame_bytes_from_env();
        ctx.abort_requested = abort_requested;
        bool wired = true;
        RingContext cross_ctx; // second level of the two-level ring
        if (use_hierarchy) {
            out.stats.plane = pcclExtPlaneTcpHierarchical;
            ctx.rank = hier.local_rank;
            ctx.world = hier.L;
            cross_ctx = ctx;
            cross_ctx.rank = hier.box;
            cross_ctx.world = hier.B;
            {
                std::lock_guard<std::mutex> lock(tx_mu_);
                if (const auto it = tx_pools_.find(hier.local_next); it != tx_pools_.end()) ctx.tx = it->second;
                if (const auto it = tx_pools_.find(hier.cross_next); it != tx_pools_.end()) cross_ctx.tx = it->second;
            }
            ctx.rx = pick_rx_pool(hier.local_prev, 10000);
            cross_ctx.rx = pick_rx_pool(hier.cross_prev, 10000);
            wired = !ctx.tx.empty() && ctx.tx.size() == ctx.rx.size() && !cross_ctx.tx.empty() && cross_ctx.tx.size() == cross_ctx.rx.size();
            if (!wired) PLOG(WARN) << "two-level ring is not fully wired for tag " << req.tag;
        } else if (world >= 2) {
            const Uuid &next = ring[(rank + 1) % world], &before = ring[(rank + world - 1) % world];
            {
                std::lock_guard<std::mutex> lock(tx_mu_);
                const auto it = tx_pools_.find(next);
                if (it != tx_pools_.end()) ctx.tx = it->second;
            }
            ctx.rx = pick_rx_pool(before, 10000);
            wired = !ctx.tx.empty() && ctx.tx.size() == ctx.rx.size();
            if (!wired) PLOG(WARN) << "no usable p2p link for tag " << req.tag;
        }
        if (wired) {
            ReduceArgs a;
            a.count = req.count;
            a.dtype = req.dtype;
            a.op = req.op;
            a.quant = req.quant;
            a.quant_dtype = req.quant_dtype;
            bool staged_ok = true;
            if (send_info.is_device || recv_info.is_device) {
                // Device buffers on the TCP plane are staged through page-locked host memory, as
                // users of the reference do by hand (python/examples/nanogpt_diloco/sync_diloco.py:494).
                // The user's buffers stay untouched until the verdict is known, so no backup copy.
                stage = staging_for(req.tag, req.count * esz);
                staged_ok = stage != nullptr;
                if (staged_ok && send_info.is_device) {
                    staged_ok = plane_->copy_to_host(stage, send, req.count * esz);
                } else if (staged_ok) {
                    std::memcpy(stage, send, req.count * esz);
                }
                a.src = a.dst = stage;
            } else {
                a.src = send;
                a.dst = recv;
            }
            if (staged_ok) {
                std::vector<uint8_t> *backup = stage || outer_two_pass || req.input_is_scratch ? nullptr : &host_backup;
                const RingResult r = use_hierarchy ? hierarchical_allreduce(ctx, cross_ctx, a, backup) : ring_allreduce(ctx, a, backup);
                local_ok = r.ok;
                out.tx_bytes = r.tx_bytes;
                out.rx_bytes = r.rx_bytes;
            }
        }
    }

    // ---- completion handshake: our verdict, then exactly one Abort packet, then Complete
    C2MCollectiveComplete done;
    done.tag = req.tag;
    done.was_aborted = !local_ok;
    if (!master_.send_packet(done)) return out;
    if (use_device && device_capable && local_ok && !aborted) {
        // The kernel's closing barrier is only passed once EVERY peer has its complete result, so
        // the outcome of the completion vote is already known.  Return now; the Abort/Complete
        // packets are consumed by the next operation on this tag or by the next joint call.
        {
            std::lock_guard<std::mutex> lock(deferred_mu_);
            deferred_[req.tag] = Deferred{abort_seen, aborted};
        }
        out.ok = apply_outer_update();
        return out;
    }
    if (!abort_seen) {
        const auto pkt = master_.recv_matching<M2CCollectiveAbort>([&](const M2CCollectiveAbort &p) { return p.tag == req.tag; });
        if (!pkt) return out;
        abort_seen = true;
        aborted = pkt->aborted;
    }
    if (!master_.recv_matching<M2CCollectiveComplete>([&](const M2CCollectiveComplete &p) { return p.tag == req.tag; })) return out;

    // A fused outer step has updated the caller's tensors once its kernel reports success (every rank held the whole
    // average at that point, so every surviving rank applied it): it stays a success even when the master aborted
    // the operation afterwards because a peer vanished; the ring is rebuilt before the next operation instead.
    const bool fused_applied = outer_fused && use_device && device_capable && local_ok;
    if (fused_applied && aborted) ring_stale_ = true;
    const bool success = fused_applied || (local_ok && !aborted);
    // (A failed device-plane attempt leaves an ordinary in-place operand as it was: the kernels write the caller's
    // memory only after every block of this rank knows the reduction completed everywhere.)
    if (out.stats.plane == pcclExtPlaneTcpRing || out.stats.plane == pcclExtPlaneTcpHierarchical) {
        if (success && stage) {
            if (recv_info.is_device) {
                if (!plane_->copy_to_device(recv, stage, req.count * esz)) return out;
            } else {
                std::memcpy(recv, stage, req.count * esz);
            }
        } else if (!success && !stage && in_place && local_ok && !host_backup.empty()) {
            // finished locally but another peer aborted: the caller will retry with this
            // buffer as input, so it must hold the original values again
            std::memcpy(recv, host_backup.data(), host_backup.size());
        }
    }
    out.ok = success && apply_outer_update();
    return out;
}

} // namespace pccl::client
