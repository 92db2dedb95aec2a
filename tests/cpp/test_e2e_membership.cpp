// Membership, topology and failure contracts (reference
// ccoip/tests/end_to_end/test_accept_new_peers.cpp, test_peer_groups.cpp, and the SIGKILL stress
// orchestrator python/tests/stress_tests/basic_stress_test/stresstest_orchestrator.py).
#include <atomic>
#include <chrono>
#include <cstdlib>
#include <random>

#include <numeric>

#include "e2e_util.hpp"
#include "pccl.h"

using namespace pccl;
using namespace std::chrono_literals;
using client::ReduceOutcome;
using client::ReduceRequest;

// connect() of the second peer blocks until the first one votes to accept it
TEST(join_blocks_until_accepted) {
    TestMaster tm;
    client::Client first(tm.config());
    CHECK(first.connect());
    CHECK_EQ(first.global_world(), 1u);
    client::Client second(tm.config(1)); // a different peer group: accept votes are global
    std::atomic<bool> joined{false};
    std::thread t([&] {
        CHECK(second.connect());
        joined = true;
    });
    std::this_thread::sleep_for(300ms);
    CHECK(!joined);
    bool pending = false;
    CHECK(first.are_peers_pending(pending));
    CHECK(pending);
    CHECK(first.update_topology());
    t.join();
    CHECK(joined);
    CHECK_EQ(first.global_world(), 2u);
    CHECK_EQ(second.global_world(), 2u);
    CHECK_EQ(first.local_world(), 1u);
    CHECK_EQ(first.num_groups(), 2u);
    CHECK_EQ(first.largest_group(), 1u);
    // nothing pending any more; the query is a joint call of all accepted peers
    std::thread q([&] {
        bool p2 = true;
        CHECK(second.are_peers_pending(p2));
        CHECK(!p2);
    });
    CHECK(first.are_peers_pending(pending));
    CHECK(!pending);
    q.join();
}

TEST(calls_before_connect_fail) {
    TestMaster tm;
    client::Client c(tm.config());
    CHECK(!c.update_topology());
    bool pending;
    CHECK(!c.are_peers_pending(pending));
    ReduceRequest r;
    float x = 1;
    r.send = r.recv = &x;
    r.count = 1;
    CHECK(c.all_reduce_async(r) == client::Client::Launch::NotConnected);
    // no master at all
    auto cfg = tm.config();
    cfg.master.port = 1;
    client::Client nowhere(cfg);
    CHECK(!nowhere.connect());
}

TEST(c_api_argument_checks) {
    CHECK_EQ(pcclInit(), pcclSuccess);
    pcclComm_t *comm = nullptr;
    CHECK_EQ(pcclCreateCommunicator(nullptr, &comm), pcclInvalidArgument);
    TestMaster tm;
    pcclCommCreateParams_t params{};
    params.master_address = tm.config().master;
    CHECK_EQ(pcclCreateCommunicator(&params, &comm), pcclSuccess);
    int v = 0;
    CHECK_EQ(pcclGetAttribute(comm, PCCL_ATTRIBUTE_GLOBAL_WORLD_SIZE, &v), pcclInvalidUsage); // not connected
    CHECK_EQ(pcclUpdateTopology(comm), pcclInvalidUsage);
    CHECK_EQ(pcclConnect(comm), pcclSuccess);
    CHECK_EQ(pcclConnect(comm), pcclInvalidUsage); // twice
    CHECK_EQ(pcclGetAttribute(comm, PCCL_ATTRIBUTE_GLOBAL_WORLD_SIZE, &v), pcclSuccess);
    CHECK_EQ(v, 1);
    CHECK_EQ(pcclGetAttribute(comm, static_cast<pcclAttribute_t>(77), &v), pcclInvalidArgument);
    float buf[4] = {1, 2, 3, 4};
    pcclReduceDescriptor_t d{};
    d.count = 4;
    d.op = pcclSum;
    d.src_descriptor.datatype = pcclFloat;
    CHECK_EQ(pcclAllReduce(buf, buf, &d, comm, nullptr), pcclTooFewPeers); // alone
    d.count = 0;
    CHECK_EQ(pcclAllReduce(buf, buf, &d, comm, nullptr), pcclInvalidArgument);
    d.count = 4;
    d.quantization_options.algorithm = pcclQuantMinMax;
    d.quantization_options.quantized_datatype = pcclInt8; // min-max needs an unsigned code type
    CHECK_EQ(pcclAllReduce(buf, buf, &d, comm, nullptr), pcclInvalidArgument);
    CHECK_EQ(pcclOptimizeTopology(comm), pcclInvalidUsage); // world of one
    CHECK_EQ(pcclExtDataTypeSize(pcclBFloat16), size_t(2));
    pcclBuildInfo_t info{};
    CHECK_EQ(pcclGetBuildInfo(&info), pcclSuccess);
    uint64_t h = 0;
    CHECK_EQ(pcclExtHashBuffer(buf, sizeof buf, pcclDeviceCpu, 0, &h), pcclSuccess);
    CHECK(h != 0);
    CHECK_EQ(pcclDestroyCommunicator(comm), pcclSuccess);
    // master handle life cycle
    pcclMasterInstance_t *m = nullptr;
    CHECK_EQ(pcclCreateMaster(tm.config().master, &m), pcclSuccess); // port is taken by `tm`
    CHECK(pcclRunMaster(m) != pcclSuccess);
    CHECK_EQ(pcclDestroyMaster(m), pcclSuccess);
}

// bandwidth measurement, ring optimisation, rewiring; reductions still work afterwards
TEST(optimize_topology_end_to_end) {
    setenv("PCCL_BENCHMARK_MILLIS", "120", 1);
    setenv("PCCL_NUM_BENCHMARK_CONNECTIONS", "2", 1);
    setenv("PCCL_BENCHMARK_BACKOFF_MILLIS", "30", 1);
    TestMaster tm;
    auto peers = connect_peers(tm, 4);
    for (int round = 0; round < 2; ++round) { // second round: every edge is known, nothing to measure
        std::vector<std::thread> ts;
        std::atomic<int> ok{0};
        for (auto &p : peers) ts.emplace_back([&] { ok += p->optimize_topology() ? 1 : 0; });
        for (auto &t : ts) t.join();
        CHECK_EQ(ok.load(), 4);
    }
    std::vector<std::thread> ts;
    for (int r = 0; r < 4; ++r)
        ts.emplace_back([&, r] {
            std::vector<double> in(10001, r + 1.0), out(10001);
            ReduceRequest q;
            q.send = in.data(), q.recv = out.data(), q.count = in.size(), q.dtype = wire::DType::F64, q.tag = 4;
            CHECK(peers[r]->all_reduce_async(q) == client::Client::Launch::Ok);
            ReduceOutcome o;
            CHECK(peers[r]->await_reduce(4, o));
            for (double v : out) CHECK(v == 10.0);
        });
    for (auto &t : ts) t.join();
}

// A peer disappears in the middle of a reduction: the survivors' operation fails with the
// input intact, the ring is rebuilt, and the retry succeeds with the smaller world.
TEST(peer_loss_during_reduce_then_retry) {
    TestMaster tm;
    auto peers = connect_peers(tm, 3);
    const size_t n = 6'000'000;
    std::atomic<bool> victim_started{false};
    std::vector<std::thread> ts;
    std::atomic<int> survivors_ok{0};
    for (int r = 0; r < 2; ++r)
        ts.emplace_back([&, r] {
            std::vector<float> buf(n, static_cast<float>(r + 1));
            int attempts = 0;
            for (;;) {
                ++attempts;
                ReduceRequest q;
                q.send = q.recv = buf.data(); // in place: a failed attempt must leave the input untouched
                q.count = n, q.op = wire::RedOp::Sum, q.tag = 1;
                CHECK(peers[r]->all_reduce_async(q) == client::Client::Launch::Ok);
                ReduceOutcome o;
                if (peers[r]->await_reduce(1, o)) {
                    // whichever world the successful attempt saw, the values must match it
                    const float want = o.world == 3 ? 6.0f : 3.0f;
                    for (size_t i = 0; i < n; i += 997) CHECK(buf[i] == want);
                    CHECK(o.world == 2 || o.world == 3);
                    break;
                }
                for (size_t i = 0; i < n; i += 997) CHECK(buf[i] == static_cast<float>(r + 1)); // rolled back
                CHECK(attempts < 10);
            }
            survivors_ok++;
        });
    std::thread victim([&] {
        std::vector<float> buf(n, 3.0f);
        ReduceRequest q;
        q.send = q.recv = buf.data();
        q.count = n, q.tag = 1;
        CHECK(peers[2]->all_reduce_async(q) == client::Client::Launch::Ok);
        victim_started = true;
        std::this_thread::sleep_for(3ms);
        peers[2]->interrupt(); // as good as a SIGKILL from the others' point of view
        ReduceOutcome o;
        peers[2]->await_reduce(1, o);
    });
    victim.join();
    for (auto &t : ts) t.join();
    CHECK_EQ(survivors_ok.load(), 2);
    // the survivors carry on as a group of two
    CHECK_EQ(peers[0]->local_world(), 2u);
    std::vector<std::thread> again;
    for (int r = 0; r < 2; ++r)
        again.emplace_back([&, r] {
            std::vector<int32_t> in(1000, r + 1), out(1000);
            ReduceRequest q;
            q.send = in.data(), q.recv = out.data(), q.count = 1000, q.dtype = wire::DType::I32, q.tag = 2;
            CHECK(peers[r]->all_reduce_async(q) == client::Client::Launch::Ok);
            ReduceOutcome o;
            CHECK(peers[r]->await_reduce(2, o));
            CHECK_EQ(o.world, 2u);
            for (auto v : out) CHECK_EQ(v, 3);
        });
    for (auto &t : again) t.join();
}

// The same on a two-level ring: a peer of a 2x2 group vanishes in the middle of a hierarchical reduction.  The survivors'
// attempt fails with the input intact; three peers cannot be split evenly over two hosts, so the retry runs on the flat ring.
TEST(peer_loss_on_a_two_level_ring) {
    TestMaster tm;
    auto peers = connect_peers(tm, 4, {}, {1, 2, 1, 2});
    for (auto &p : peers) CHECK(p->hierarchy_shape() == std::make_pair(2u, 2u));
    const size_t n = 6'000'000;
    std::vector<std::thread> ts;
    std::atomic<int> survivors_ok{0};
    for (int r = 0; r < 3; ++r)
        ts.emplace_back([&, r] {
            std::vector<float> buf(n, static_cast<float>(r + 1));
            int attempts = 0;
            for (;;) {
                ++attempts;
                ReduceRequest q;
                q.send = q.recv = buf.data();
                q.count = n, q.op = wire::RedOp::Sum, q.tag = 1;
                CHECK(peers[r]->all_reduce_async(q) == client::Client::Launch::Ok);
                ReduceOutcome o;
                if (peers[r]->await_reduce(1, o)) {
                    const float want = o.world == 4 ? 10.0f : 6.0f;
                    for (size_t i = 0; i < n; i += 997) CHECK(buf[i] == want);
                    CHECK(o.world == 3 || o.world == 4);
                    CHECK_EQ(static_cast<int>(o.stats.plane), static_cast<int>(o.world == 4 ? pcclExtPlaneTcpHierarchical : pcclExtPlaneTcpRing));
                    break;
                }
                for (size_t i = 0; i < n; i += 997) CHECK(buf[i] == static_cast<float>(r + 1)); // rolled back
                CHECK(attempts < 10);
            }
            survivors_ok++;
        });
    std::thread victim([&] {
        std::vector<float> buf(n, 4.0f);
        ReduceRequest q;
        q.send = q.recv = buf.data();
        q.count = n, q.tag = 1;
        CHECK(peers[3]->all_reduce_async(q) == client::Client::Launch::Ok);
        std::this_thread::sleep_for(3ms);
        peers[3]->interrupt();
        ReduceOutcome o;
        peers[3]->await_reduce(1, o);
    });
    victim.join();
    for (auto &t : ts) t.join();
    CHECK_EQ(survivors_ok.load(), 3);
    CHECK_EQ(peers[0]->local_world(), 3u);
    for (int r = 0; r < 3; ++r) CHECK(peers[r]->hierarchy_shape() == std::make_pair(0u, 0u));
}

// Two dozen peers in one run, three groups of eight: admission one by one (every admission is a vote of everybody already
// in), ring per group, concurrent reductions in all groups, then a third of the peers leave at once and the rest carry on.
TEST(two_dozen_peers_three_groups) {
    TestMaster tm;
    const int n_peers = std::getenv("PCCL_TEST_MANY_PEERS") ? 3 * (std::atoi(std::getenv("PCCL_TEST_MANY_PEERS")) / 3) : 24;
    const uint32_t per_group = static_cast<uint32_t>(n_peers / 3);
    std::vector<uint32_t> groups(n_peers);
    for (int i = 0; i < n_peers; ++i) groups[i] = static_cast<uint32_t>(i % 3);
    auto peers = connect_peers(tm, n_peers, groups);
    CHECK_EQ(peers[0]->global_world(), uint64_t(n_peers));
    CHECK_EQ(peers[0]->local_world(), per_group);
    CHECK_EQ(peers[0]->num_groups(), 3u);
    auto reduce_all = [&](const std::vector<int> &who, uint32_t want_world, uint64_t tag) {
        std::vector<std::thread> ts;
        std::atomic<int> good{0};
        for (int r : who)
            ts.emplace_back([&, r] {
                std::vector<int32_t> in(50'000, r % 3 + 1), out(50'000);
                ReduceRequest q;
                q.send = in.data(), q.recv = out.data(), q.count = in.size(), q.dtype = wire::DType::I32, q.op = wire::RedOp::Sum, q.tag = tag;
                ReduceOutcome o;
                if (peers[r]->all_reduce_sync(q, o) != client::Client::Launch::Ok || !o.ok) return;
                if (o.world == want_world && out[0] == static_cast<int32_t>(want_world) * (r % 3 + 1) && out.back() == out[0]) good++;
            });
        for (auto &t : ts) t.join();
        return good.load();
    };
    std::vector<int> everybody(n_peers);
    std::iota(everybody.begin(), everybody.end(), 0);
    CHECK_EQ(reduce_all(everybody, per_group, 1), n_peers);
    // peers 0..8 vanish (three from every group); the others notice at their next joint call
    for (int r = 0; r < 9; ++r) peers[r]->interrupt();
    std::vector<int> rest;
    for (int r = 9; r < n_peers; ++r) rest.push_back(r);
    bool settled = false;
    for (int attempt = 0; attempt < 20 && !settled; ++attempt) {
        std::vector<std::thread> ts;
        for (int r : rest) ts.emplace_back([&, r] { peers[r]->update_topology(); });
        for (auto &t : ts) t.join();
        settled = peers[9]->global_world() == uint64_t(n_peers - 9) && peers[9]->local_world() == per_group - 3;
    }
    CHECK(settled);
    CHECK_EQ(reduce_all(rest, per_group - 3, 2), n_peers - 9);
}

// pcclAllReduceMultipleWithRetry keeps going when a peer vanishes
TEST(multiple_with_retry_survives_peer_loss) {
    CHECK_EQ(pcclInit(), pcclSuccess);
    TestMaster tm;
    pcclCommCreateParams_t params{};
    params.master_address = tm.config().master;
    constexpr int W = 3, kOps = 12;
    constexpr size_t n = 300'000;
    pcclComm_t *comms[W] = {};
    // peers join one after the other, the earlier ones admit the later ones
    for (int i = 0; i < W; ++i) {
        CHECK_EQ(pcclCreateCommunicator(&params, &comms[i]), pcclSuccess);
        std::atomic<bool> joined{false};
        std::thread j([&] {
            CHECK_EQ(pcclConnect(comms[i]), pcclSuccess);
            joined = true;
        });
        while (!joined) {
            int world = 0;
            if (i > 0 && pcclGetAttribute(comms[0], PCCL_ATTRIBUTE_GLOBAL_WORLD_SIZE, &world) == pcclSuccess && world < i + 1) {
                std::vector<std::thread> ts;
                for (int k = 0; k < i; ++k) ts.emplace_back([&, k] { CHECK_EQ(pcclUpdateTopology(comms[k]), pcclSuccess); });
                for (auto &t : ts) t.join();
            } else {
                std::this_thread::sleep_for(2ms);
            }
        }
        j.join();
    }
    std::vector<std::thread> ts;
    std::atomic<int> done{0};
    for (int r = 0; r < W; ++r)
        ts.emplace_back([&, r] {
            std::vector<std::vector<float>> in(kOps, std::vector<float>(n, 1.0f)), out(kOps, std::vector<float>(n, 0.0f));
            std::vector<pcclReduceOpDescriptor_t> descs(kOps);
            for (int k = 0; k < kOps; ++k) {
                descs[k].sendbuf = in[k].data();
                descs[k].recvbuf = out[k].data();
                descs[k].descriptor.count = n;
                descs[k].descriptor.op = pcclSum;
                descs[k].descriptor.tag = 100 + k;
                descs[k].descriptor.src_descriptor.datatype = pcclFloat;
                descs[k].descriptor.quantization_options.algorithm = pcclQuantNone;
            }
            if (r == 2) { // the victim starts and dies
                pcclAsyncReduceOp_t h{};
                pcclAllReduceAsync(in[0].data(), out[0].data(), &descs[0].descriptor, comms[r], &h);
                std::this_thread::sleep_for(2ms);
                pcclDestroyCommunicator(comms[r]);
                comms[r] = nullptr;
                return;
            }
            pcclReduceInfo_t info{};
            CHECK_EQ(pcclAllReduceMultipleWithRetry(descs.data(), kOps, comms[r], &info, 4), pcclSuccess);
            for (int k = 0; k < kOps; ++k) {
                const float v = out[k][n / 2];
                CHECK(v == 2.0f || v == 3.0f); // done either before or after the loss
            }
            CHECK(info.tx_bytes > 0);
            done++;
        });
    for (auto &t : ts) t.join();
    CHECK_EQ(done.load(), 2);
    for (auto *c : comms)
        if (c) CHECK_EQ(pcclDestroyCommunicator(c), pcclSuccess);
}

TEST(master_loss_fails_cleanly) {
    auto tm = std::make_unique<TestMaster>();
    auto peers = connect_peers(*tm, 2);
    tm.reset(); // master gone
    std::this_thread::sleep_for(100ms);
    for (auto &p : peers) {
        CHECK(!p->update_topology());
        ReduceRequest q;
        float x = 1, y = 0;
        q.send = &x, q.recv = &y, q.count = 1;
        const auto l = p->all_reduce_async(q);
        if (l == client::Client::Launch::Ok) {
            ReduceOutcome o;
            CHECK(!p->await_reduce(0, o));
        }
    }
}

// The master is the one process every peer depends on: junk on its port (port scanners, peers of another version,
// truncated or absurd frames, frames in the wrong state) may cost the sender its connection and nothing else.
TEST(master_survives_garbage) {
    TestMaster tm;
    std::mt19937_64 rng(20240607);
    auto target = tm.config().master;
    auto random_bytes = [&](size_t n) {
        std::vector<uint8_t> v(n);
        for (auto &b : v) b = static_cast<uint8_t>(rng());
        return v;
    };
    auto frame = [&](uint16_t id, const std::vector<uint8_t> &payload) {
        std::vector<uint8_t> f;
        const uint64_t len = 2 + payload.size();
        for (int i = 7; i >= 0; --i) f.push_back(static_cast<uint8_t>(len >> (8 * i)));
        f.push_back(static_cast<uint8_t>(id >> 8));
        f.push_back(static_cast<uint8_t>(id));
        f.insert(f.end(), payload.begin(), payload.end());
        return f;
    };
    for (int round = 0; round < 200; ++round) {
        net::BlockingSocket sock(target);
        CHECK(sock.connect());
        const int kind = static_cast<int>(rng() % 5);
        if (kind == 0) { // no framing at all
            const auto junk = random_bytes(1 + rng() % 300);
            sock.send_raw(junk.data(), junk.size());
        } else if (kind == 1) { // a length prefix that promises more than ever arrives
            auto f = frame(static_cast<uint16_t>(rng() % 14), random_bytes(rng() % 64));
            f[6] = 0x40;
            sock.send_raw(f.data(), f.size());
        } else {
            if (kind >= 3) { // become a registered (the first: accepted) client, then misbehave
                wire::C2MSessionRegistration reg;
                reg.peer_group = static_cast<uint32_t>(rng() % 3);
                reg.p2p_port = 1000;
                reg.ss_port = 1001;
                reg.bm_port = 1002;
                sock.send_packet(reg);
            }
            const int n = 1 + static_cast<int>(rng() % 6);
            for (int i = 0; i < n; ++i) {
                // every known id with a random body, among them list counts of 2^63 and strings longer than the frame
                auto body = random_bytes(rng() % 120);
                if (!body.empty() && rng() % 3 == 0) std::fill(body.begin(), body.begin() + std::min<size_t>(8, body.size()), 0x7f);
                const auto f = frame(static_cast<uint16_t>(rng() % 14), body);
                if (!sock.send_raw(f.data(), f.size())) break;
            }
        }
        if (rng() % 2) std::this_thread::sleep_for(std::chrono::milliseconds(1));
    } // sockets close here
    std::this_thread::sleep_for(200ms);
    // the master still does its job
    auto peers = connect_peers(tm, 2);
    CHECK_EQ(peers[0]->global_world(), 2u);
    std::vector<float> a(1000, 1.0f), b(1000, 2.0f), ra(1000), rb(1000);
    std::thread t([&] {
        ReduceRequest req;
        req.send = b.data();
        req.recv = rb.data();
        req.count = 1000;
        req.tag = 1;
        ReduceOutcome out;
        CHECK(peers[1]->all_reduce_sync(req, out) == client::Client::Launch::Ok && out.ok);
    });
    ReduceRequest req;
    req.send = a.data();
    req.recv = ra.data();
    req.count = 1000;
    req.tag = 1;
    ReduceOutcome out;
    CHECK(peers[0]->all_reduce_sync(req, out) == client::Client::Launch::Ok && out.ok);
    t.join();
    CHECK(ra[0] == 3.0f && rb[999] == 3.0f);
}

// The same for a peer's three listening ports (ring neighbours, shared-state requests, bandwidth probes).
TEST(peer_listeners_survive_garbage) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::mt19937_64 rng(77);
    for (int round = 0; round < 150; ++round) {
        auto target = tm.config().master; // 127.0.0.1, port replaced below
        client::Client &victim = *peers[round % 2];
        const uint16_t ports[3] = {victim.p2p_port(), victim.shared_state_port(), victim.benchmark_port()};
        target.port = ports[rng() % 3];
        net::BlockingSocket sock(target);
        if (!sock.connect()) continue;
        std::vector<uint8_t> junk(1 + rng() % 400);
        for (auto &b : junk) b = static_cast<uint8_t>(rng());
        if (junk.size() > 8 && rng() % 2) { // plausible framing: small length, then noise
            for (int i = 0; i < 7; ++i) junk[i] = 0;
            junk[7] = static_cast<uint8_t>(junk.size() - 8);
        }
        sock.send_raw(junk.data(), junk.size());
    }
    std::this_thread::sleep_for(200ms);
    // both peers still reduce and synchronise
    for (int tag = 0; tag < 3; ++tag) {
        std::vector<float> a(5000, 1.0f), b(5000, 2.0f);
        std::thread t([&] {
            ReduceRequest req;
            req.send = req.recv = b.data();
            req.count = b.size();
            req.tag = static_cast<uint64_t>(tag);
            ReduceOutcome out;
            CHECK(peers[1]->all_reduce_sync(req, out) == client::Client::Launch::Ok && out.ok);
        });
        ReduceRequest req;
        req.send = req.recv = a.data();
        req.count = a.size();
        req.tag = static_cast<uint64_t>(tag);
        ReduceOutcome out;
        CHECK(peers[0]->all_reduce_sync(req, out) == client::Client::Launch::Ok && out.ok);
        t.join();
        CHECK(a[0] == 3.0f && b[4999] == 3.0f);
    }
}

TEST_MAIN()
