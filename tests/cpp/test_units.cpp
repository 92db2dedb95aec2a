// Unit tests of the host-side building blocks: wire codec (golden bytes hand-derived from the
// reference serialisers, ccoip/src/cpp/ccoip_packets.cpp), ATSP solver, bandwidth store, hashes,
// quantisers, 16-bit float conversions.
#include "device/sequencer.hpp"
#include "kernels/mset_formula.hpp"
#include <sys/mman.h>
#include <random>
#include <algorithm>
#include <thread>
#include <algorithm>
#include <chrono>
#include <typeinfo>
#include <cmath>
#include <cstring>
#include <numeric>
#include <random>
#include <thread>

#include <sys/wait.h>
#include <unistd.h>

#include "harness.hpp"
#include "kernels/cpu/half.hpp"
#include "kernels/cpu/hash.hpp"
#include "kernels/cpu/quantize.hpp"
#include "kernels/cpu/quantize_fast.hpp"
#include "kernels/cpu/reduce.hpp"
#include "topo/atsp.hpp"
#include "topo/bandwidth_store.hpp"
#include "topo/optimizer.hpp"
#include "util/guard.hpp"
#include "util/guarded_alloc.hpp"
#include "util/park.hpp"
#include "util/ring_queue.hpp"
#include "wire/packets.hpp"

using namespace pccl;
using namespace pccl::wire;

static std::vector<uint8_t> bytes_of(std::initializer_list<int> v) {
    std::vector<uint8_t> out;
    for (int x : v) out.push_back(static_cast<uint8_t>(x));
    return out;
}

TEST(wire_primitives_are_big_endian) {
    Writer w;
    w.put<uint16_t>(0x1234);
    w.put<uint32_t>(0xdeadbeef);
    w.put<uint64_t>(1);
    w.put<bool>(true);
    w.put<float>(1.0f);
    w.put_string("ab");
    CHECK(w.bytes() == bytes_of({0x12, 0x34, 0xde, 0xad, 0xbe, 0xef, 0, 0, 0, 0, 0, 0, 0, 1, 1, 0x3f, 0x80, 0, 0, 0, 0, 0, 0, 0, 0, 0, 2, 'a', 'b'}));
    Reader r(w.bytes().data(), w.size());
    CHECK_EQ(r.get<uint16_t>(), 0x1234);
    CHECK_EQ(r.get<uint32_t>(), 0xdeadbeefu);
    CHECK_EQ(r.get<uint64_t>(), 1u);
    CHECK(r.get<bool>());
    CHECK(r.get<float>() == 1.0f);
    CHECK(r.get_string() == "ab");
    CHECK(r.ok() && r.remaining() == 0);
    r.get<uint32_t>(); // reading past the end is an error, not a crash
    CHECK(!r.ok());
}

TEST(wire_session_registration_golden) {
    C2MSessionRegistration p;
    p.peer_group = 7;
    p.p2p_port = 48149, p.ss_port = 48150, p.bm_port = 48151;
    const auto frame = encode_frame(p);
    // [u64 len = 2 + 11][u16 id = 1][u32 group][bool explicit = 0][3 x u16 port]
    CHECK(frame == bytes_of({0, 0, 0, 0, 0, 0, 0, 13, 0, 1, 0, 0, 0, 7, 0, 0xbc, 0x15, 0xbc, 0x16, 0xbc, 0x17}));
    C2MSessionRegistration q;
    Reader r(frame.data() + 10, frame.size() - 10);
    CHECK(q.read(r));
    CHECK_EQ(q.peer_group, 7u);
    CHECK_EQ(q.ss_port, 48150);
    // explicit addresses: three socket addresses instead of ports
    p.explicit_addresses = true;
    p.adv_p2p.inet.protocol = inetIPv4;
    p.adv_p2p.inet.ipv4 = {{10, 1, 2, 3}};
    p.adv_p2p.port = 5000;
    p.adv_ss = p.adv_bm = p.adv_p2p;
    Writer w;
    p.write(w);
    CHECK_EQ(w.size(), size_t(4 + 1 + 3 * (1 + 4 + 2)));
    CHECK(std::vector<uint8_t>(w.bytes().begin() + 5, w.bytes().begin() + 12) == bytes_of({1, 10, 1, 2, 3, 0x13, 0x88}));
}

TEST(wire_collective_packets_golden) {
    C2MCollectiveInitiate p;
    p.tag = 0x0102030405060708ull;
    p.count = 4;
    p.dtype = DType::F32;
    p.op = RedOp::Sum;
    Writer w;
    p.write(w);
    CHECK(w.bytes() == bytes_of({1, 2, 3, 4, 5, 6, 7, 8, 0, 0, 0, 0, 0, 0, 0, 4, 10, 1}));
    M2CCollectiveCommence c;
    c.tag = 5, c.seq_nr = 9;
    Writer w2;
    c.write(w2);
    CHECK(w2.bytes() == bytes_of({0, 0, 0, 0, 0, 0, 0, 5, 0, 0, 0, 0, 0, 0, 0, 9}));
    CHECK_EQ(C2MCollectiveInitiate::id, 10);
    CHECK_EQ(C2MCollectiveComplete::id, 11);
    CHECK_EQ(M2CCollectiveCommence::id, 9);
    CHECK_EQ(M2CCollectiveAbort::id, 11);
}

TEST(wire_round_trips) {
    std::mt19937_64 rng(1);
    M2CP2PConnectionInfo info;
    info.global_world_size = 8, info.local_world_size = 4, info.num_peer_groups = 2, info.largest_group_size = 4;
    for (int i = 0; i < 3; ++i) {
        PeerEndpoint e;
        e.uuid = Uuid::random_v4();
        e.p2p_addr.inet.protocol = i == 1 ? inetIPv6 : inetIPv4;
        e.p2p_addr.port = static_cast<uint16_t>(1000 + i);
        info.peers.push_back(e);
    }
    Writer w;
    info.write(w);
    M2CP2PConnectionInfo back;
    Reader r(w.bytes().data(), w.size());
    CHECK(back.read(r) && r.remaining() == 0);
    CHECK_EQ(back.peers.size(), size_t(3));
    CHECK(back.peers[1].uuid == info.peers[1].uuid && back.peers[1].p2p_addr.inet.protocol == inetIPv6);
    info.unchanged = true; // the peer list is not on the wire then
    Writer w2;
    info.write(w2);
    CHECK_EQ(w2.size(), size_t(1 + 4 * 8));

    C2MSyncSharedState s;
    s.revision = 12;
    s.strategy = SyncStrategy::TxOnly;
    s.entries.push_back({"layer.0.weight", 0xabcdef, HashType::Simple, 1024, DType::BF16, false});
    s.entries.push_back({"step", 0, HashType::Simple, 1, DType::I64, true});
    Writer w3;
    s.write(w3);
    C2MSyncSharedState s2;
    Reader r3(w3.bytes().data(), w3.size());
    CHECK(s2.read(r3) && s2.entries == s.entries && s2.revision == 12 && s2.strategy == SyncStrategy::TxOnly);

    // truncated input must be rejected, never read out of bounds
    for (size_t cut = 0; cut + 1 < w3.size(); cut += 3) {
        C2MSyncSharedState bad;
        Reader rb(w3.bytes().data(), cut);
        CHECK(!bad.read(rb));
    }
    // absurd element counts are refused before allocating
    Writer evil;
    evil.put<uint64_t>(1), evil.put<uint8_t>(0), evil.put<uint64_t>(~0ull);
    C2MSyncSharedState bad;
    Reader re(evil.bytes().data(), evil.size());
    CHECK(!bad.read(re));

    P2PDequantMeta m;
    m.tag = 3;
    m.meta_type = DequantMetaType::ZeroPointScale;
    m.dtype_a = DType::I64, m.blob_a = bytes_of({0, 0, 0, 0, 0, 0, 0, 9});
    m.dtype_b = DType::F32, m.blob_b = bytes_of({1, 2, 3, 4});
    Writer w4;
    m.write(w4);
    P2PDequantMeta m2;
    Reader r4(w4.bytes().data(), w4.size());
    CHECK(m2.read(r4) && m2.blob_a == m.blob_a && m2.blob_b == m.blob_b && m2.dtype_b == DType::F32);
}

// Every packet decoder against hostile input: random bytes of random lengths, and valid encodings cut short at every
// length.  A decoder may accept or refuse, but it never reads out of bounds (the sanitizer builds run this too), never
// allocates according to a count it has not checked against the bytes that are there, and never takes long.
template <typename P> static void fuzz_decoder(std::mt19937_64 &rng, const P *valid = nullptr) {
    std::vector<uint8_t> junk(300);
    for (int round = 0; round < 400; ++round) {
        const size_t n = rng() % junk.size();
        for (size_t i = 0; i < n; ++i) junk[i] = static_cast<uint8_t>(round % 3 == 0 ? 0xff : rng());
        P p;
        Reader r(junk.data(), n);
        (void) p.read(r);
    }
    if (valid) {
        Writer w;
        valid->write(w);
        P whole;
        Reader rw(w.bytes().data(), w.size());
        const bool whole_ok = whole.read(rw) && rw.remaining() == 0;
        if (!whole_ok) std::printf("decoder of %s refuses its own encoding (%zu bytes, %zu left)\n", typeid(P).name(), w.size(), rw.remaining());
        CHECK(whole_ok);
        for (size_t cut = 0; cut < w.size(); ++cut) {
            P p;
            Reader r(w.bytes().data(), cut);
            const bool refused = !p.read(r);
            if (!refused) std::printf("decoder of %s accepts %zu of %zu bytes\n", typeid(P).name(), cut, w.size());
            CHECK(refused);
        }
    }
}

TEST(packet_decoders_survive_hostile_input) {
    std::mt19937_64 rng(2024);
    const auto t0 = std::chrono::steady_clock::now();
    C2MSessionRegistration reg;
    reg.peer_group = 3;
    fuzz_decoder(rng, &reg);
    C2MP2PEstablished est;
    est.success = false;
    est.failed_peers = {Uuid::random_v4(), Uuid::random_v4()};
    fuzz_decoder(rng, &est);
    C2MReportBandwidth bw;
    fuzz_decoder(rng, &bw);
    C2MSyncSharedState sync;
    sync.entries.push_back({"w", 7, HashType::Mset64, 10, DType::F32, false});
    fuzz_decoder(rng, &sync);
    C2MCollectiveInitiate init;
    init.tag = 9, init.count = 1000;
    fuzz_decoder(rng, &init);
    C2MCollectiveComplete done;
    fuzz_decoder(rng, &done);
    fuzz_decoder<C2MEstablishP2P>(rng);
    fuzz_decoder<M2CSessionRegistrationResponse>(rng);
    M2CP2PConnectionInfo info;
    info.peers.resize(2);
    fuzz_decoder(rng, &info);
    M2CP2PEstablished ring;
    ring.success = true;
    ring.ring_order = {Uuid::random_v4(), Uuid::random_v4(), Uuid::random_v4()};
    fuzz_decoder(rng, &ring);
    fuzz_decoder<M2CPeersPendingResponse>(rng);
    M2COptimizeTopologyResponse opt;
    fuzz_decoder(rng, &opt);
    fuzz_decoder<M2COptimizeTopologyComplete>(rng);
    M2CSyncSharedState verdict;
    verdict.is_outdated = true;
    verdict.outdated_keys = {"a", "bb"};
    verdict.expected_hashes = {1, 2};
    verdict.expected_hash_types = {HashType::Simple, HashType::Crc32};
    fuzz_decoder(rng, &verdict);
    fuzz_decoder<M2CCollectiveCommence>(rng);
    fuzz_decoder<M2CCollectiveComplete>(rng);
    fuzz_decoder<M2CCollectiveAbort>(rng);
    M2CLeaseRevoked revoked;
    revoked.tag = 77, revoked.all = true, revoked.peer_lost = true;
    fuzz_decoder(rng, &revoked);
    {
        Writer w;
        revoked.write(w);
        M2CLeaseRevoked back;
        Reader r(w.bytes().data(), w.size());
        CHECK(back.read(r) && back.tag == 77 && back.all && back.peer_lost);
    }
    P2PHello hello;
    fuzz_decoder(rng, &hello);
    P2PDequantMeta meta;
    meta.blob_a = {1, 2, 3, 4}, meta.blob_b = {5, 6, 7, 8};
    fuzz_decoder(rng, &meta);
    P2PDirectory dir;
    for (int i = 0; i < 3; ++i) {
        DirectoryEntry e;
        e.uuid = Uuid::random_v4();
        e.host_id = 1000 + i, e.has_host = i != 1, e.has_endpoint = i != 0;
        e.endpoint.inet.protocol = i == 2 ? inetIPv6 : inetIPv4;
        e.endpoint.port = static_cast<uint16_t>(40000 + i);
        dir.entries.push_back(e);
    }
    fuzz_decoder(rng, &dir);
    {
        Writer w;
        dir.write(w);
        P2PDirectory back;
        Reader r(w.bytes().data(), w.size());
        CHECK(back.read(r) && back.entries.size() == 3);
        for (int i = 0; i < 3; ++i)
            CHECK(back.entries[i].uuid == dir.entries[i].uuid && back.entries[i].host_id == dir.entries[i].host_id && back.entries[i].has_host == dir.entries[i].has_host &&
                  back.entries[i].has_endpoint == dir.entries[i].has_endpoint && back.entries[i].endpoint.port == dir.entries[i].endpoint.port);
        Writer evil; // a count that the bytes cannot back
        evil.put<uint64_t>(4000);
        P2PDirectory bad;
        Reader re(evil.bytes().data(), evil.size());
        CHECK(!bad.read(re));
        CHECK(bad.entries.size() < 4000);
    }
    C2SRequestSharedState want;
    want.keys = {"x", "yy", "zzz"};
    fuzz_decoder(rng, &want);
    S2CSharedStateResponse have;
    have.entries.push_back({"x", 123});
    fuzz_decoder(rng, &have);
    C2SRequestSharedStateP2P want2;
    want2.keys = {"k"};
    fuzz_decoder(rng, &want2);
    S2CSharedStateResponseP2P have2;
    have2.entries.push_back({"k", 5});
    fuzz_decoder(rng, &have2);
    fuzz_decoder<C2BHello>(rng);
    fuzz_decoder<B2CBusy>(rng);
    CHECK(std::chrono::steady_clock::now() - t0 < std::chrono::seconds(20));
}

static double brute_force(const std::vector<double> &cost, int n) {
    std::vector<int> perm(n - 1);
    std::iota(perm.begin(), perm.end(), 1);
    double best = topo::kNoEdge;
    do {
        std::vector<int> tour{0};
        tour.insert(tour.end(), perm.begin(), perm.end());
        best = std::min(best, topo::atsp_tour_cost(cost, n, tour));
    } while (std::next_permutation(perm.begin(), perm.end()));
    return best;
}

TEST(atsp_exact_matches_brute_force) {
    std::mt19937 rng(7);
    std::uniform_real_distribution<double> d(0.1, 10.0);
    for (int n = 2; n <= 8; ++n)
        for (int rep = 0; rep < 5; ++rep) {
            std::vector<double> cost(n * n);
            for (auto &c : cost) c = d(rng);
            if (rep == 4 && n > 3) cost[1 * n + 2] = cost[2 * n + 1] = topo::kNoEdge; // missing link
            const auto r = topo::atsp_solve_exact(cost, n);
            const double want = brute_force(cost, n);
            CHECK(r.feasible && r.optimal);
            CHECK(std::fabs(r.cost - want) < 1e-9);
            CHECK(std::fabs(topo::atsp_tour_cost(cost, n, r.tour) - r.cost) < 1e-9);
            std::vector<int> sorted = r.tour;
            std::sort(sorted.begin(), sorted.end());
            for (int i = 0; i < n; ++i) CHECK_EQ(sorted[i], i);
        }
}

TEST(atsp_heuristic_is_close_and_respects_missing_edges) {
    std::mt19937 rng(11);
    std::uniform_real_distribution<double> d(0.1, 10.0);
    for (int n : {9, 12, 24}) {
        std::vector<double> cost(n * n);
        for (auto &c : cost) c = d(rng);
        const auto h = topo::atsp_solve_heuristic(cost, n, 500, 4 * n, 3);
        CHECK(h.feasible);
        if (n <= 12) {
            const auto e = topo::atsp_solve_exact(cost, n);
            CHECK(h.cost <= e.cost * 1.25 + 1e-9);
        }
    }
    // only one Hamiltonian cycle exists: 0->1->2->3->4->0
    const int n = 5;
    std::vector<double> cost(n * n, topo::kNoEdge);
    for (int i = 0; i < n; ++i) cost[i * n + (i + 1) % n] = 1.0;
    const auto h = topo::atsp_solve_heuristic(cost, n, 200, 20, 5);
    CHECK(h.feasible && std::fabs(h.cost - 5.0) < 1e-9);
}

TEST(bandwidth_store_and_ring_optimiser) {
    topo::BandwidthStore store;
    std::vector<Uuid> peers;
    for (int i = 0; i < 4; ++i) {
        peers.push_back(Uuid::random_v4());
        CHECK(store.register_peer(peers.back()));
    }
    CHECK(!store.register_peer(peers[0]));
    CHECK_EQ(store.missing_edges(peers[0]).size(), size_t(6));
    CHECK(!store.fully_populated());
    // fast ring 0->2->1->3->0, everything else slow
    auto fast = [&](int a, int b) { return (a == 0 && b == 2) || (a == 2 && b == 1) || (a == 1 && b == 3) || (a == 3 && b == 0); };
    for (int a = 0; a < 4; ++a)
        for (int b = 0; b < 4; ++b)
            if (a != b) CHECK(store.store(peers[a], peers[b], fast(a, b) ? 40000.0 : 100.0));
    CHECK(store.fully_populated());
    CHECK(!store.store(peers[0], peers[0], 1.0));
    std::vector<Uuid> ring = peers;
    bool optimal = false;
    CHECK(topo::optimize_ring(store, ring, optimal));
    CHECK(optimal);
    // the ring must follow the fast edges
    for (size_t i = 0; i < ring.size(); ++i) {
        const int a = static_cast<int>(std::find(peers.begin(), peers.end(), ring[i]) - peers.begin());
        const int b = static_cast<int>(std::find(peers.begin(), peers.end(), ring[(i + 1) % 4]) - peers.begin());
        CHECK(fast(a, b));
    }
    CHECK(store.unregister_peer(peers[3]));
    CHECK(!store.get(peers[0], peers[3]));
    CHECK(store.debug_matrix().find("Gbit/s") != std::string::npos);
}

static uint32_t lcg(uint32_t &s) {
    s = 1664525u * s + 1013904223u;
    return s % 256u;
}

TEST(simplehash_golden_values) {
    // reference golden values: ccoip/tests/unit_tests/simple_hash/simplehash_cuda_test.cpp:58,104,150
    for (auto [n, want] : {std::pair<size_t, uint64_t>{1, 344386053ull}, {4, 3765247898ull}, {25, 3651434421ull}}) {
        std::vector<uint8_t> d(n);
        uint32_t s = 42;
        for (auto &b : d) b = static_cast<uint8_t>(lcg(s));
        CHECK_EQ(cpu::simplehash(d.data(), n), want);
    }
    CHECK_EQ(cpu::simplehash(nullptr, 0), 0u);
    // unaligned host pointers are fine
    std::vector<uint8_t> d(4099);
    uint32_t s = 1;
    for (auto &b : d) b = static_cast<uint8_t>(lcg(s));
    std::vector<uint8_t> shifted(d.size() + 1);
    std::memcpy(shifted.data() + 1, d.data(), d.size());
    CHECK_EQ(cpu::simplehash(d.data(), d.size()), cpu::simplehash(shifted.data() + 1, d.size()));
}

TEST(simplehash_golden_147MiB) {
    // reference golden value: simplehash_cuda_test.cpp:195 (154 533 888 bytes of the seed-42 LCG)
    const size_t n = 154533888;
    std::vector<uint8_t> d(n);
    uint32_t s = 42;
    for (auto &b : d) b = static_cast<uint8_t>(lcg(s));
    CHECK_EQ(cpu::simplehash(d.data(), n), 3391090508ull);
}

TEST(crc32c_all_paths_agree) {
    CHECK_EQ(cpu::crc32c("123456789", 9), 0xE3069283u); // standard check value
    std::vector<uint8_t> d(8u << 20);
    uint32_t s = 3;
    for (auto &b : d) b = static_cast<uint8_t>(lcg(s));
    for (size_t n : {size_t(0), size_t(1), size_t(7), size_t(8), size_t(255), size_t(256 * 3), size_t(8192 * 3 + 5), d.size() - 3, d.size()}) {
        for (size_t off : {size_t(0), size_t(1), size_t(5)}) {
            if (off + n > d.size()) continue;
            const uint32_t hw = cpu::crc32c(d.data() + off, n);
            cpu::crc32c_force_generic(true); // what the reference calls "spoofing" CPU features
            const uint32_t sw = cpu::crc32c(d.data() + off, n);
            cpu::crc32c_force_generic(false);
            CHECK_EQ(hw, sw);
            CHECK_EQ(sw, cpu::crc32c_generic(d.data() + off, n));
        }
    }
    const uint32_t a = cpu::crc32c(d.data(), 1000), b = cpu::crc32c(d.data() + 1000, 5000);
    CHECK_EQ(cpu::crc32c_combine(a, b, 5000), cpu::crc32c(d.data(), 6000));
}

TEST(half_conversions) {
    for (uint32_t h = 0; h < 65536; ++h) {
        const float f = cpu::f16_to_f32(static_cast<uint16_t>(h));
        if (std::isnan(f)) continue;
        CHECK_EQ(cpu::f32_to_f16(f), static_cast<uint16_t>(h));
        const float g = cpu::bf16_to_f32(static_cast<uint16_t>(h));
        if (std::isnan(g)) continue;
        CHECK_EQ(cpu::f32_to_bf16(g), static_cast<uint16_t>(h));
    }
    CHECK_EQ(cpu::f32_to_f16(65520.0f), 0x7c00); // rounds to infinity
    CHECK_EQ(cpu::f32_to_f16(1.0f + 1.0f / 2048.0f), 0x3c00); // tie to even
    CHECK_EQ(cpu::f32_to_bf16(1.0f + 1.0f / 256.0f), 0x3f80);
}

TEST(minmax_quantiser_matches_reference_formula) {
    std::vector<float> x(1000);
    for (size_t i = 0; i < x.size(); ++i) x[i] = -3.0f + 0.007f * static_cast<float>(i); // also covers all-negative sub-ranges
    cpu::QuantParams p;
    p.algo = QuantAlgo::MinMax;
    p.code_dtype = DType::U8;
    cpu::quant_analyze(x.data(), x.size(), p);
    CHECK(p.min == x.front() && p.max == x.back());
    std::vector<uint8_t> q(x.size());
    cpu::quantize(x.data(), q.data(), x.size(), p);
    for (size_t i = 0; i < x.size(); ++i) {
        const double v = (double(x[i]) - double(x.front())) * (1.0 / double(x.back() - x.front())) * 255.0;
        CHECK_EQ(q[i], static_cast<uint8_t>(v)); // truncating, in double
    }
    std::vector<float> y(x.size(), 0.0f);
    cpu::dequantize_reduce(y.data(), q.data(), x.size(), p, RedOp::Set);
    for (size_t i = 0; i < x.size(); ++i) CHECK(std::fabs(y[i] - x[i]) <= (x.back() - x.front()) / 255.0f + 1e-6f);
    // all-negative chunk keeps a sane maximum (the reference seeds its max scan with FLT_MIN)
    std::vector<float> neg{-5.0f, -4.0f, -4.5f};
    cpu::quant_analyze(neg.data(), neg.size(), p);
    CHECK(p.max == -4.0f && p.min == -5.0f);
    // metadata survives the wire
    cpu::QuantParams back = p;
    back.min = back.max = 0;
    CHECK(cpu::from_wire(cpu::to_wire(p, 1), back));
    CHECK(back.min == p.min && back.max == p.max);
}

// The vectorised float32 / 8-bit paths against the generic templates that define the arithmetic: identical codes, identical
// metadata, identical decoded values -- on ordinary data, on ragged lengths, and on the values where rounding decides.
TEST(fast_quantisers_equal_the_generic_ones_bit_for_bit) {
    std::mt19937_64 rng(99);
    std::vector<std::vector<float>> inputs;
    for (size_t n : {size_t(0), size_t(1), size_t(3), size_t(127), size_t(128), size_t(129), size_t(1000), size_t(65536 + 77)}) {
        std::vector<float> x(n);
        std::normal_distribution<float> normal(0.5f, 3.0f);
        for (auto &v : x) v = normal(rng);
        inputs.push_back(x);
    }
    { // values of very different magnitudes, exact ties, zeros of both signs, subnormals, huge values, a constant block
        std::vector<float> x;
        for (int e = -30; e <= 30; ++e)
            for (float m : {1.0f, 1.0625f, 1.125f, 1.1875f, 1.5f, 1.9375f, 1.99999f}) x.push_back(std::ldexp(m, e)), x.push_back(-std::ldexp(m, e));
        for (float v : {0.0f, -0.0f, 1e-40f, -1e-42f, 3.4e38f, -3.4e38f, 448.0f, 447.9f, 464.0f, 0.0009765625f, 0.00146484375f}) x.push_back(v);
        x.resize(((x.size() + 127) / 128 + 1) * 128, 7.25f); // last block constant
        inputs.push_back(x);
        std::vector<float> same(300, -2.5f);
        inputs.push_back(same);
        std::vector<float> broken = inputs[6]; // a gradient that has gone wrong in places: same garbage out of both
        broken[5] = std::numeric_limits<float>::quiet_NaN(), broken[200] = std::numeric_limits<float>::infinity(),
        broken[700] = -std::numeric_limits<float>::infinity(), broken[701] = std::numeric_limits<float>::quiet_NaN();
        inputs.push_back(broken);
    }
    struct Case {
        QuantAlgo algo;
        DType code;
    };
    for (const Case c : {Case{QuantAlgo::MinMax, DType::U8}, Case{QuantAlgo::ZeroPointScale, DType::U8}, Case{QuantAlgo::BlockMinMax, DType::U8},
                         Case{QuantAlgo::BlockFp8, DType::F8E4M3}})
        for (const auto &x : inputs) {
            cpu::QuantParams pf, pg;
            pf.algo = pg.algo = c.algo;
            pf.code_dtype = pg.code_dtype = c.code;
            std::vector<uint8_t> qf(cpu::quantized_size(pf, x.size()) + 1, 0xAB), qg(qf.size(), 0xAB);
            std::vector<float> acc(x.size());
            for (size_t i = 0; i < acc.size(); ++i) acc[i] = 0.1f * static_cast<float>(i % 17) - 0.5f;
            std::vector<float> set_f(x.size(), 1.0f), set_g(x.size(), 1.0f), add_f = acc, add_g = acc, avg_f = acc, avg_g = acc;
            cpu::quant_force_generic(false);
            cpu::quant_analyze(x.data(), x.size(), pf);
            cpu::quantize(x.data(), qf.data(), x.size(), pf);
            cpu::dequantize_reduce(set_f.data(), qf.data(), x.size(), pf, RedOp::Set);
            cpu::dequantize_reduce(add_f.data(), qf.data(), x.size(), pf, RedOp::Sum);
            cpu::dequantize_reduce(avg_f.data(), qf.data(), x.size(), pf, RedOp::Avg);
            cpu::quant_force_generic(true);
            cpu::quant_analyze(x.data(), x.size(), pg);
            cpu::quantize(x.data(), qg.data(), x.size(), pg);
            cpu::dequantize_reduce(set_g.data(), qg.data(), x.size(), pg, RedOp::Set);
            cpu::dequantize_reduce(add_g.data(), qg.data(), x.size(), pg, RedOp::Sum);
            cpu::dequantize_reduce(avg_g.data(), qg.data(), x.size(), pg, RedOp::Avg);
            cpu::quant_force_generic(false);
            CHECK(pf.min == pg.min && pf.max == pg.max && pf.scale == pg.scale && pf.zero_point == pg.zero_point);
            CHECK(qf == qg);
            CHECK_EQ(qf.back(), uint8_t(0xAB)); // nothing written past the end
            auto same_bits = [&](const std::vector<float> &a, const std::vector<float> &b) { return x.empty() || std::memcmp(a.data(), b.data(), x.size() * 4) == 0; };
            CHECK(same_bits(set_f, set_g));
            CHECK(same_bits(add_f, add_g));
            CHECK(same_bits(avg_f, avg_g));
        }
    // the fp8 conversions themselves: every code, and a few hundred million floats spread over all exponents (every 13th
    // bit pattern would take a minute; a stride of 4099 covers each binade with thousands of mantissas, ties included)
    for (int c = 0; c < 256; ++c) {
        const float a = cpu::fp8e4m3_to_f32(static_cast<uint8_t>(c)), b = cpu::fast::fp8e4m3_to_f32_branchless(static_cast<uint8_t>(c));
        CHECK(std::memcmp(&a, &b, 4) == 0);
    }
    uint64_t mismatches = 0;
    for (uint64_t u = 0; u <= 0xffffffffull; u += 4099) {
        const float f = std::bit_cast<float>(static_cast<uint32_t>(u));
        mismatches += cpu::f32_to_fp8e4m3(f) != cpu::fast::f32_to_fp8e4m3_branchless(f);
    }
    for (int c = 0; c < 256; ++c) // exact code points and the midpoints between neighbours
        for (int d = 0; d < 2; ++d) {
            const float lo = cpu::fp8e4m3_to_f32(static_cast<uint8_t>(c)), hi = cpu::fp8e4m3_to_f32(static_cast<uint8_t>((c & 0x7f) == 0x7e ? c : c + 1));
            const float f = d ? 0.5f * (lo + hi) : lo;
            if (f == f) mismatches += cpu::f32_to_fp8e4m3(f) != cpu::fast::f32_to_fp8e4m3_branchless(f);
        }
    CHECK_EQ(mismatches, uint64_t(0));
}

// The 16-bit float loops (F16C instructions, branch-free bfloat16 rounding) against the conversions of half.hpp that define
// them: every operator, the averaging division, every bit pattern on one side and awkward partners on the other.
TEST(half_precision_reductions_equal_their_definition) {
    std::vector<uint16_t> all(65536 + 5); // + a ragged tail
    for (size_t i = 0; i < all.size(); ++i) all[i] = static_cast<uint16_t>(i);
    std::mt19937 rng(5);
    std::vector<uint16_t> partner(all.size());
    for (auto &v : partner) v = static_cast<uint16_t>(rng());
    const uint16_t special[] = {0x0000, 0x8000, 0x0001, 0x83ff, 0x0400, 0x3c00, 0xbc00, 0x7bff, 0xfbff, 0x7c00, 0xfc00, 0x7e00, 0x7c01, 0x7f80, 0xff80, 0x7fc0, 0x3f80};
    for (size_t i = 0; i < partner.size(); i += 97) partner[i] = special[(i / 97) % (sizeof special / 2)];
    struct Fmt {
        DType dtype;
        float (*load)(uint16_t);
        uint16_t (*store)(float);
    };
    for (const Fmt f : {Fmt{DType::F16, cpu::f16_to_f32, cpu::f32_to_f16}, Fmt{DType::BF16, cpu::bf16_to_f32, cpu::f32_to_bf16}}) {
        for (RedOp op : {RedOp::Sum, RedOp::Prod, RedOp::Max, RedOp::Min, RedOp::Set}) {
            std::vector<uint16_t> got = all;
            cpu::reduce_inplace(got.data(), partner.data(), got.size(), f.dtype, op);
            size_t bad = 0;
            for (size_t i = 0; i < all.size(); ++i) {
                const float a = f.load(all[i]), b = f.load(partner[i]);
                uint16_t want;
                switch (op) {
                case RedOp::Sum: want = f.store(a + b); break;
                case RedOp::Prod: want = f.store(a * b); break;
                case RedOp::Max: want = f.store(a > b ? a : b); break;
                case RedOp::Min: want = f.store(a < b ? a : b); break;
                default: want = partner[i]; break;
                }
                // two NaN operands: which of the two payloads (and signs) survives an x86 addition depends on the order the
                // compiler hands them to the instruction, in the definition as much as in the fast loops; any NaN will do
                if (a != a && b != b && op != RedOp::Set) want = f.store(f.load(got[i])) == got[i] && f.load(got[i]) != f.load(got[i]) ? got[i] : want;
                bad += got[i] != want;
            }
            CHECK_EQ(bad, size_t(0));
        }
        for (uint64_t world : {uint64_t(2), uint64_t(3), uint64_t(7)}) {
            std::vector<uint16_t> got = all;
            cpu::finalize_avg(got.data(), got.size(), f.dtype, world);
            size_t bad = 0;
            for (size_t i = 0; i < all.size(); ++i) bad += got[i] != f.store(f.load(all[i]) / static_cast<float>(world));
            CHECK_EQ(bad, size_t(0));
        }
    }
}

// The averaging division on a buffer large enough to be cut into ranges for several threads: every element divided exactly
// once, ragged length, the two types whose division is not a plain float divide included.
TEST(average_of_a_large_buffer_is_divided_once_everywhere) {
    const size_t n = (48u << 20) / 4 + 13;
    std::vector<float> f(n);
    std::vector<int32_t> k(n);
    for (size_t i = 0; i < n; ++i) f[i] = static_cast<float>(i % 1000) * 3.0f, k[i] = static_cast<int32_t>(i % 1000) * 7 - 3000;
    cpu::finalize_avg(f.data(), n, DType::F32, 3);
    cpu::finalize_avg(k.data(), n, DType::I32, 7);
    size_t bad = 0;
    for (size_t i = 0; i < n; ++i) {
        bad += f[i] != static_cast<float>(i % 1000) * 3.0f / 3.0f;
        bad += k[i] != (static_cast<int32_t>(i % 1000) * 7 - 3000) / 7;
    }
    CHECK_EQ(bad, size_t(0));
    std::vector<uint16_t> h((40u << 20) / 2 + 5, cpu::f32_to_bf16(6.0f));
    cpu::finalize_avg(h.data(), h.size(), DType::BF16, 4);
    for (size_t i = 0; i < h.size(); i += 4099) CHECK_EQ(h[i], cpu::f32_to_bf16(1.5f));
    CHECK_EQ(h.back(), cpu::f32_to_bf16(1.5f));
}

TEST(zero_point_scale_quantiser) {
    std::vector<float> x(512);
    for (size_t i = 0; i < x.size(); ++i) x[i] = std::sin(0.05f * static_cast<float>(i)) * 4.0f + 1.0f;
    for (DType code : {DType::U8, DType::I8, DType::U16, DType::I16}) {
        cpu::QuantParams p;
        p.algo = QuantAlgo::ZeroPointScale;
        p.code_dtype = code;
        cpu::quant_analyze(x.data(), x.size(), p);
        std::vector<uint8_t> q(cpu::quantized_size(p, x.size()));
        cpu::quantize(x.data(), q.data(), x.size(), p);
        std::vector<float> y(x.size(), 0.0f);
        cpu::dequantize_reduce(y.data(), q.data(), x.size(), p, RedOp::Set);
        const float step = static_cast<float>(p.scale);
        for (size_t i = 0; i < x.size(); ++i) CHECK(std::fabs(y[i] - x[i]) <= 0.51f * step + 1e-6f);
        cpu::QuantParams back = p;
        back.scale = 0, back.zero_point = 0;
        CHECK(cpu::from_wire(cpu::to_wire(p, 1), back));
        CHECK(back.zero_point == p.zero_point && std::fabs(back.scale - p.scale) < 1e-12);
    }
    // constants survive exactly (the contract behind "42 + 43 == 85")
    for (float c : {42.0f, -7.25f, 0.0f}) {
        std::vector<float> k(100, c), y(100, 1.0f);
        cpu::QuantParams p;
        p.algo = QuantAlgo::ZeroPointScale;
        p.code_dtype = DType::U8;
        cpu::quant_analyze(k.data(), k.size(), p);
        std::vector<uint8_t> q(k.size());
        cpu::quantize(k.data(), q.data(), k.size(), p);
        cpu::dequantize_reduce(y.data(), q.data(), k.size(), p, RedOp::Sum);
        for (float v : y) CHECK(v == c + 1.0f);
    }
}

TEST(reduce_kernels_all_types) {
    std::vector<int16_t> a{1, -2, 30000}, b{5, 7, 2767};
    cpu::reduce_inplace(a.data(), b.data(), 3, DType::I16, RedOp::Sum);
    CHECK(a[0] == 6 && a[1] == 5 && a[2] == static_cast<int16_t>(32767));
    std::vector<uint64_t> u{10, 20}, v{3, 40};
    cpu::reduce_inplace(u.data(), v.data(), 2, DType::U64, RedOp::Max);
    CHECK(u[0] == 10 && u[1] == 40);
    cpu::reduce_inplace(u.data(), v.data(), 2, DType::U64, RedOp::Prod);
    CHECK(u[0] == 30 && u[1] == 1600);
    std::vector<int32_t> avg{7, -7, 9};
    cpu::finalize_avg(avg.data(), 3, DType::I32, 2);
    CHECK(avg[0] == 3 && avg[1] == -3 && avg[2] == 4); // integer division truncates
    std::vector<uint16_t> h{cpu::f32_to_bf16(1.5f)}, g{cpu::f32_to_bf16(2.25f)};
    cpu::reduce_inplace(h.data(), g.data(), 1, DType::BF16, RedOp::Sum);
    CHECK(cpu::bf16_to_f32(h[0]) == 3.75f);
    CHECK_EQ(wire::dtype_size(DType::BF16), size_t(2)); // the reference reports 0 here (src/pccl.cpp:525-543)
    CHECK_EQ(wire::dtype_size(DType::I16), size_t(2));
    CHECK_EQ(wire::dtype_size(DType::F16), size_t(2));
}

TEST(mpsc_ring_keeps_per_producer_order) {
    MpscRing<std::pair<int, int>> ring(64);
    CHECK_EQ(ring.capacity(), 64u);
    // full / empty behaviour
    for (int i = 0; i < 64; ++i) CHECK(ring.try_push({0, i}));
    std::pair<int, int> extra{9, 9}, got;
    CHECK(!ring.try_push(std::move(extra)));
    for (int i = 0; i < 64; ++i) CHECK(ring.try_pop(got) && got.second == i);
    CHECK(!ring.try_pop(got));
    // four producers, one consumer: nothing lost, nothing duplicated, each producer's items arrive in order
    constexpr int kProducers = 4, kItems = 20000;
    std::vector<std::thread> producers;
    for (int p = 0; p < kProducers; ++p)
        producers.emplace_back([&ring, p] {
            for (int i = 0; i < kItems; ++i) {
                std::pair<int, int> item{p, i};
                while (!ring.try_push(std::move(item))) std::this_thread::yield();
            }
        });
    int next[kProducers] = {0, 0, 0, 0};
    for (int received = 0; received < kProducers * kItems;) {
        if (!ring.try_pop(got)) {
            std::this_thread::yield();
            continue;
        }
        CHECK_EQ(got.second, next[got.first]);
        ++next[got.first];
        ++received;
    }
    for (auto &t : producers) t.join();
    CHECK(!ring.try_pop(got));
    // elements left inside are destroyed with the ring
    auto counter = std::make_shared<int>(0);
    {
        MpscRing<std::shared_ptr<int>> owned(8);
        for (int i = 0; i < 5; ++i) CHECK(owned.try_push(std::shared_ptr<int>(counter)));
        CHECK_EQ(counter.use_count(), 6);
    }
    CHECK_EQ(counter.use_count(), 1);
}

TEST(parker_wakes_and_times_out) {
    Parker p(0);
    const auto t0 = std::chrono::steady_clock::now();
    CHECK_EQ(p.wait_while(0, 30), 0u); // nobody publishes: times out with the old value
    const auto waited = std::chrono::duration_cast<std::chrono::milliseconds>(std::chrono::steady_clock::now() - t0).count();
    CHECK(waited >= 25 && waited < 2000);
    for (int round = 0; round < 200; ++round) {
        Parker q(0);
        std::thread waker([&] { q.publish(7); });
        CHECK_EQ(q.wait_while(0, 5000), 7u);
        waker.join();
    }
    CHECK_EQ(p.wait_while(1, -1), 0u); // value already differs: returns at once even without a limit
}

TEST(phase_guard_is_exclusive) {
    std::atomic<bool> flag{false};
    {
        PhaseGuard outer(flag);
        CHECK(outer.entered());
        PhaseGuard inner(flag);
        CHECK(!inner.entered());
        CHECK(flag.load());
    }
    CHECK(!flag.load());
}

TEST(thread_guard_adopts_first_thread) {
    ThreadGuard g;
    std::thread([&] { g.check("unit"); }).join();
    g.release();
    g.check("unit"); // adopted by this thread now
    g.check("unit");
    // a foreign thread is a fatal error: observe it in a child process
    const pid_t pid = fork();
    if (pid == 0) {
        std::thread([&] { g.check("unit"); }).join();
        _exit(0);
    }
    int status = 0;
    waitpid(pid, &status, 0);
    CHECK(!(WIFEXITED(status) && WEXITSTATUS(status) == 0));
}

// ---- same-box sequencer: processes that ask for the same operations in different orders get ONE common order
TEST(sequencer_gives_every_process_the_same_order) {
    using pccl::device::Sequencer;
    void *mem = mmap(nullptr, Sequencer::bytes(), PROT_READ | PROT_WRITE, MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    CHECK(mem != MAP_FAILED);
    Sequencer::format(mem, Sequencer::fresh_base());
    constexpr int kProcs = 4, kTags = 5, kOps = 200;
    int pipes[kProcs][2];
    std::vector<pid_t> kids;
    for (int p = 0; p < kProcs; ++p) {
        CHECK(pipe(pipes[p]) == 0);
        const pid_t pid = fork();
        if (pid == 0) {
            Sequencer seq;
            if (!seq.attach(mem)) _exit(2);
            // every process issues the same (tag, k) set, each tag in order, tags interleaved differently per process
            std::vector<uint64_t> next(kTags, 0), got(kTags * kOps);
            std::mt19937 rng(1234 + p);
            for (int done = 0; done < kTags * kOps;) {
                const int t = static_cast<int>(rng() % kTags);
                if (next[t] == kOps) continue;
                uint64_t index = 0;
                if (!seq.order_of(100 + t, next[t], index)) _exit(3);
                got[t * kOps + next[t]] = index;
                ++next[t];
                ++done;
            }
            const ssize_t n = write(pipes[p][1], got.data(), got.size() * sizeof(uint64_t));
            _exit(n == static_cast<ssize_t>(got.size() * sizeof(uint64_t)) ? 0 : 4);
        }
        kids.push_back(pid);
    }
    std::vector<std::vector<uint64_t>> all(kProcs, std::vector<uint64_t>(kTags * kOps));
    for (int p = 0; p < kProcs; ++p) {
        size_t off = 0;
        while (off < all[p].size() * sizeof(uint64_t)) {
            const ssize_t n = read(pipes[p][0], reinterpret_cast<uint8_t *>(all[p].data()) + off, all[p].size() * sizeof(uint64_t) - off);
            CHECK(n > 0);
            off += static_cast<size_t>(n);
        }
        int status = 0;
        waitpid(kids[p], &status, 0);
        CHECK(WIFEXITED(status) && WEXITSTATUS(status) == 0);
    }
    for (int p = 1; p < kProcs; ++p) CHECK(all[p] == all[0]); // same index for the same operation everywhere
    std::vector<uint64_t> sorted = all[0];
    std::sort(sorted.begin(), sorted.end());
    for (size_t i = 0; i < sorted.size(); ++i) CHECK_EQ(sorted[i], i); // a permutation of 0..n-1: no gaps, no duplicates
    for (int t = 0; t < kTags; ++t)
        for (int k = 1; k < kOps; ++k) CHECK(all[0][t * kOps + k] > all[0][t * kOps + k - 1]); // per tag the order is the issue order
    Sequencer a, b;
    CHECK(a.attach(mem));
    CHECK(a.flag_epoch(5) == a.base() + 5);
    std::this_thread::sleep_for(std::chrono::milliseconds(3));
    CHECK(Sequencer::fresh_base() > a.flag_epoch(kTags * kOps)); // a later log starts above everything this one handed out
    munmap(mem, Sequencer::bytes());
}

TEST(mset64_is_order_independent_and_position_dependent) {
    std::vector<uint8_t> data(1 << 20);
    std::mt19937 rng(7);
    for (auto &b : data) b = static_cast<uint8_t>(rng());
    const uint64_t whole = pccl::cpu::mset64(data.data(), data.size());
    // any split into spans of whole packets gives the same raw sum
    uint64_t raw = 0;
    size_t off = 0;
    for (size_t len : {16u * 1000u, 16u * 1u, 16u * 40000u}) {
        raw += pccl::cpu::mset64_raw(data.data() + off, len, off / 16);
        off += len;
    }
    raw += pccl::cpu::mset64_raw(data.data() + off, data.size() - off, off / 16);
    CHECK_EQ(pccl::mset::finalize(raw, data.size()), whole);
    // a flipped bit, swapped packets and a different length all change the value
    data[12345] ^= 1;
    CHECK(pccl::cpu::mset64(data.data(), data.size()) != whole);
    data[12345] ^= 1;
    for (int i = 0; i < 16; ++i) std::swap(data[i], data[16 + i]);
    CHECK(pccl::cpu::mset64(data.data(), data.size()) != whole);
    for (int i = 0; i < 16; ++i) std::swap(data[i], data[16 + i]);
    CHECK_EQ(pccl::cpu::mset64(data.data(), data.size()), whole);
    CHECK(pccl::cpu::mset64(data.data(), data.size() - 1) != whole);
    // tails: zero padding is not confused with real zeros because the length is part of the value
    const uint8_t three[3] = {1, 2, 3}, four[4] = {1, 2, 3, 0};
    CHECK(pccl::cpu::mset64(three, 3) != pccl::cpu::mset64(four, 4));
    CHECK_EQ(pccl::cpu::mset64(nullptr, 0), pccl::mset::finalize(0, 0));
}

TEST(guarded_blocks_fault_on_overrun) {
    for (size_t n : {1u, 15u, 16u, 4096u, 4097u, 100000u}) {
        auto *p = static_cast<uint8_t *>(guarded_alloc(n));
        CHECK(p != nullptr);
        CHECK_EQ(reinterpret_cast<uintptr_t>(p) % 16, 0u);
        std::memset(p, 0x11, n);
        CHECK(guarded_free(p));
    }
    CHECK_EQ(guarded_live_blocks(), 0u);
    // one byte past a block whose size is a multiple of 16 must fault; so must the byte before a page-sized block
    for (int which = 0; which < 2; ++which) {
        const pid_t pid = fork();
        if (pid == 0) {
            auto *p = static_cast<volatile uint8_t *>(guarded_alloc(4096));
            if (which == 0) p[4096] = 1;
            else p[-1] = 1;
            _exit(0);
        }
        int status = 0;
        waitpid(pid, &status, 0);
        // a plain build dies of SIGSEGV; a sanitizer build catches the fault and exits non-zero
        CHECK(WIFSIGNALED(status) ? WTERMSIG(status) == SIGSEGV : WEXITSTATUS(status) != 0);
    }
    // an underrun into the alignment slack is caught by the canary
    auto *p = static_cast<uint8_t *>(guarded_alloc(40));
    p[-1] = 0;
    CHECK(!guarded_free(p));
}

TEST_MAIN()
