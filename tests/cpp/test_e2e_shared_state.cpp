// Shared-state synchronisation contracts (reference
// ccoip/tests/end_to_end/test_shared_state_distribution.cpp:45-3073).
#include <cstring>
#include <numeric>

#include "e2e_util.hpp"

using namespace pccl;
using client::SharedEntry;
using client::SharedStateArgs;
using wire::SyncStrategy;

struct PeerState {
    std::vector<std::vector<uint8_t>> buffers;
    SharedStateArgs args;
    bool ok = false;
    uint64_t tx = 0, rx = 0;
    void add(const std::string &key, size_t n, uint8_t fill, wire::DType dtype = wire::DType::U8, bool allow_ineq = false) {
        buffers.emplace_back(n * wire::dtype_size(dtype), fill);
        SharedEntry e;
        e.key = key;
        e.count = n;
        e.dtype = dtype;
        e.allow_content_inequality = allow_ineq;
        args.entries.push_back(e);
        for (size_t i = 0; i < buffers.size(); ++i) args.entries[i].data = buffers[i].data();
    }
};

static void sync_all(std::vector<std::unique_ptr<client::Client>> &peers, std::vector<PeerState> &st, const std::vector<int> &who = {}) {
    std::vector<std::thread> ts;
    for (size_t i = 0; i < peers.size(); ++i) {
        if (!who.empty() && std::find(who.begin(), who.end(), static_cast<int>(i)) == who.end()) continue;
        ts.emplace_back([&, i] { st[i].ok = peers[i]->sync_shared_state(st[i].args, st[i].tx, st[i].rx); });
    }
    for (auto &t : ts) t.join();
}

// two peers, different content: exactly one direction transfers `size` bytes; afterwards equal
TEST(basic_two_peers_differ) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::vector<PeerState> st(2);
    st[0].add("w", 1024, 7);
    st[1].add("w", 1024, 9);
    sync_all(peers, st);
    CHECK(st[0].ok && st[1].ok);
    CHECK(st[0].buffers[0] == st[1].buffers[0]);
    CHECK_EQ(st[0].tx + st[1].tx, 1024u);
    CHECK_EQ(st[0].rx + st[1].rx, 1024u);
    CHECK((st[0].tx == 1024 && st[1].rx == 1024) || (st[1].tx == 1024 && st[0].rx == 1024));
}

TEST(identical_content_moves_nothing) {
    TestMaster tm;
    auto peers = connect_peers(tm, 3);
    std::vector<PeerState> st(3);
    for (auto &s : st) s.add("w", 4096, 3, wire::DType::F32);
    sync_all(peers, st);
    for (auto &s : st) CHECK(s.ok && s.tx == 0 && s.rx == 0);
}

// only keys whose hash differs are transferred
TEST(only_dirty_keys_move) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::vector<PeerState> st(2);
    for (int p = 0; p < 2; ++p) {
        st[p].add("a", 1000, 1);
        st[p].add("b", 2000, p == 0 ? 2 : 5);
        st[p].add("c", 3000, 3, wire::DType::BF16);
    }
    sync_all(peers, st);
    CHECK(st[0].ok && st[1].ok);
    CHECK_EQ(st[0].tx + st[1].tx, 2000u);
    CHECK(st[0].buffers[1] == st[1].buffers[1]);
}

// popularity beats arrival order: 2 of 3 peers agree, the odd one receives
TEST(popular_content_wins) {
    TestMaster tm;
    auto peers = connect_peers(tm, 3);
    std::vector<PeerState> st(3);
    st[0].add("w", 512, 42);
    st[1].add("w", 512, 11);
    st[2].add("w", 512, 11);
    // the minority peer votes first
    std::thread first([&] { st[0].ok = peers[0]->sync_shared_state(st[0].args, st[0].tx, st[0].rx); });
    std::this_thread::sleep_for(std::chrono::milliseconds(100));
    sync_all(peers, st, {1, 2});
    first.join();
    CHECK(st[0].ok && st[1].ok && st[2].ok);
    CHECK_EQ(st[0].rx, 512u);
    CHECK(st[0].buffers[0] == std::vector<uint8_t>(512, 11));
    CHECK_EQ(st[1].rx + st[2].rx, 0u);
}

// peer groups are isolated: different content and key sets across groups, no traffic, no kicks
TEST(peer_groups_are_isolated) {
    TestMaster tm;
    auto peers = connect_peers(tm, 4, {0, 0, 1, 1});
    std::vector<PeerState> st(4);
    st[0].add("w", 256, 1);
    st[1].add("w", 256, 1);
    st[2].add("other", 100, 2);
    st[2].add("w", 256, 9); // same key name as in group 0, different content: no conflict
    st[3].add("other", 100, 2);
    st[3].add("w", 256, 9);
    sync_all(peers, st);
    for (auto &s : st) CHECK(s.ok && s.tx == 0 && s.rx == 0);
    CHECK_EQ(peers[0]->local_world(), 2u);
    CHECK_EQ(peers[0]->global_world(), 4u);
    CHECK_EQ(peers[0]->num_groups(), 2u);
}

// multi-step: everybody advances content identically and bumps the revision by one
TEST(multi_step_revisions) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::vector<PeerState> st(2);
    for (auto &s : st) s.add("w", 2048, 0);
    for (int step = 0; step < 5; ++step) {
        for (auto &s : st) {
            std::fill(s.buffers[0].begin(), s.buffers[0].end(), static_cast<uint8_t>(step));
            s.args.revision = 10 + step; // the first revision of a group may be anything (resume)
        }
        sync_all(peers, st);
        for (auto &s : st) CHECK(s.ok && s.tx == 0 && s.rx == 0);
    }
}

// drag-along: a peer that never bumps its revision is outdated; it adopts the distributor's
// revision even when no bytes move
TEST(outdated_revision_is_dragged_along) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::vector<PeerState> st(2);
    for (auto &s : st) s.add("w", 1024, 5);
    st[0].args.revision = st[1].args.revision = 1;
    sync_all(peers, st);
    CHECK(st[0].ok && st[1].ok);
    st[0].args.revision = 2; // leader moves on, laggard stays at 1 with equal content
    sync_all(peers, st);
    CHECK(st[0].ok && st[1].ok);
    CHECK_EQ(st[1].args.revision, 2u);
    CHECK_EQ(st[1].rx, 0u);
    // now the leader also changes content: the laggard receives it
    st[0].args.revision = 3;
    std::fill(st[0].buffers[0].begin(), st[0].buffers[0].end(), uint8_t(77));
    sync_all(peers, st);
    CHECK(st[0].ok && st[1].ok);
    CHECK_EQ(st[1].args.revision, 3u);
    CHECK_EQ(st[1].rx, 1024u);
    CHECK(st[1].buffers[0] == st[0].buffers[0]);
}

// revision may only grow by one: the violator is kicked; the lone outdated peer then fails too
TEST(one_increment_rule) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::vector<PeerState> st(2);
    for (auto &s : st) s.add("w", 64, 1);
    st[0].args.revision = st[1].args.revision = 1;
    sync_all(peers, st);
    CHECK(st[0].ok && st[1].ok);
    st[0].args.revision = 4; // expected 2
    st[1].args.revision = 1; // outdated
    sync_all(peers, st);
    CHECK(!st[0].ok);
    CHECK(!st[1].ok);
}

// key set / dtype mismatch against the elected mask: the offender is kicked, the others go on
TEST(key_set_mismatch_is_kicked) {
    TestMaster tm;
    auto peers = connect_peers(tm, 3);
    std::vector<PeerState> st(3);
    st[0].add("w", 128, 1);
    st[1].add("w", 128, 1);
    st[2].add("different_key", 128, 1);
    sync_all(peers, st);
    CHECK(st[0].ok && st[1].ok);
    CHECK(!st[2].ok);
    // survivors keep working
    st[0].args.revision = st[1].args.revision = 1;
    sync_all(peers, st, {0, 1});
    CHECK(st[0].ok && st[1].ok);
    // the kicked peer's later calls fail
    CHECK(!peers[2]->sync_shared_state(st[2].args, st[2].tx, st[2].rx));
}

TEST(send_only_and_receive_only) {
    TestMaster tm;
    auto peers = connect_peers(tm, 3);
    std::vector<PeerState> st(3);
    st[0].add("w", 300, 8);
    st[1].add("w", 300, 1);
    st[2].add("w", 300, 2);
    st[0].args.strategy = SyncStrategy::TxOnly;
    st[1].args.strategy = st[2].args.strategy = SyncStrategy::RxOnly;
    sync_all(peers, st);
    for (auto &s : st) CHECK(s.ok);
    CHECK_EQ(st[0].tx, 600u);
    CHECK_EQ(st[0].rx, 0u);
    CHECK(st[1].buffers[0] == st[0].buffers[0] && st[2].buffers[0] == st[0].buffers[0]);
}

// both SEND_ONLY with different content: at least one is kicked
TEST(conflicting_send_only) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::vector<PeerState> st(2);
    st[0].add("w", 64, 1);
    st[1].add("w", 64, 2);
    st[0].args.strategy = st[1].args.strategy = SyncStrategy::TxOnly;
    sync_all(peers, st);
    CHECK(!st[0].ok || !st[1].ok);
}

// everybody RECEIVE_ONLY: nobody offers content, the whole group is kicked
TEST(all_receive_only_kicks_group) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::vector<PeerState> st(2);
    st[0].add("w", 64, 1);
    st[1].add("w", 64, 1);
    st[0].args.strategy = st[1].args.strategy = SyncStrategy::RxOnly;
    sync_all(peers, st);
    CHECK(!st[0].ok && !st[1].ok);
}

// ENFORCE_POPULAR may not be mixed with other strategies: the others are kicked, the enforcer succeeds
TEST(enforce_popular_cannot_be_mixed) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::vector<PeerState> st(2);
    st[0].add("w", 64, 1);
    st[1].add("w", 64, 1);
    st[0].args.strategy = SyncStrategy::EnforcePopular;
    st[1].args.strategy = SyncStrategy::TxOnly;
    sync_all(peers, st);
    CHECK(st[0].ok);
    CHECK(!st[1].ok);
}

// ENFORCE_POPULAR x RECEIVE_ONLY (reference test_shared_state_distribution.cpp:2766): the receiver is kicked, the enforcer succeeds
TEST(enforce_popular_with_receive_only) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::vector<PeerState> st(2);
    st[0].add("w", 64, 1);
    st[1].add("w", 64, 2);
    st[0].args.strategy = SyncStrategy::EnforcePopular;
    st[1].args.strategy = SyncStrategy::RxOnly;
    sync_all(peers, st);
    CHECK(st[0].ok);
    CHECK(!st[1].ok);
    CHECK_EQ(st[0].tx, 0u);
}

// three peers, two enforcers and one SEND_ONLY (reference :2848): only the odd one out is kicked, the enforcers agree
TEST(enforce_popular_three_peers_one_send_only) {
    TestMaster tm;
    auto peers = connect_peers(tm, 3);
    std::vector<PeerState> st(3);
    st[0].add("w", 128, 5);
    st[1].add("w", 128, 6); // unpopular among the enforcers
    st[2].add("w", 128, 5);
    st[0].args.strategy = st[1].args.strategy = SyncStrategy::EnforcePopular;
    st[2].args.strategy = SyncStrategy::TxOnly;
    sync_all(peers, st);
    CHECK(st[0].ok && st[1].ok);
    CHECK(!st[2].ok);
    CHECK(st[0].buffers[0] == st[1].buffers[0]); // the survivors hold one state
}

// three peers, one enforcer among a SEND_ONLY and a RECEIVE_ONLY peer (reference :2960): both others are kicked
TEST(enforce_popular_three_peers_mixed_others) {
    TestMaster tm;
    auto peers = connect_peers(tm, 3);
    std::vector<PeerState> st(3);
    for (int i = 0; i < 3; ++i) st[i].add("w", 32, static_cast<uint8_t>(10 + i));
    st[0].args.strategy = SyncStrategy::EnforcePopular;
    st[1].args.strategy = SyncStrategy::TxOnly;
    st[2].args.strategy = SyncStrategy::RxOnly;
    sync_all(peers, st);
    CHECK(st[0].ok);
    CHECK(!st[1].ok && !st[2].ok);
    // the enforcer keeps working alone afterwards... with nobody left there is nothing to agree on, the call still succeeds
    st[0].args.revision = 1;
    uint64_t tx = 0, rx = 0;
    CHECK(peers[0]->sync_shared_state(st[0].args, tx, rx));
    CHECK_EQ(tx + rx, 0u);
}

// The outcome of an election does not depend on the order in which the votes reach the master: the same three peers
// (an enforcer, a SEND_ONLY peer and a RECEIVE_ONLY peer, all with different content) vote in every order, and two
// enforcers with different content always settle on the same one of the two.
TEST(elections_do_not_depend_on_vote_order) {
    const int orders[6][3] = {{0, 1, 2}, {0, 2, 1}, {1, 0, 2}, {1, 2, 0}, {2, 0, 1}, {2, 1, 0}};
    for (const auto &order : orders) {
        TestMaster tm;
        auto peers = connect_peers(tm, 3);
        std::vector<PeerState> st(3);
        for (int i = 0; i < 3; ++i) st[i].add("w", 32, static_cast<uint8_t>(10 + i));
        st[0].args.strategy = SyncStrategy::EnforcePopular;
        st[1].args.strategy = SyncStrategy::TxOnly;
        st[2].args.strategy = SyncStrategy::RxOnly;
        std::vector<std::thread> ts;
        for (int k = 0; k < 3; ++k) {
            const int p = order[k];
            ts.emplace_back([&, p] { st[p].ok = peers[p]->sync_shared_state(st[p].args, st[p].tx, st[p].rx); });
            std::this_thread::sleep_for(std::chrono::milliseconds(60));
        }
        for (auto &t : ts) t.join();
        CHECK(st[0].ok);
        CHECK(!st[1].ok && !st[2].ok);
        CHECK(st[0].buffers[0] == std::vector<uint8_t>(32, 10)); // the enforcer kept its own content
    }
    std::vector<uint8_t> settled;
    for (int first = 0; first < 2; ++first) {
        TestMaster tm;
        auto peers = connect_peers(tm, 2);
        std::vector<PeerState> st(2);
        st[0].add("w", 48, 77);
        st[1].add("w", 48, 99);
        std::thread early([&] { st[first].ok = peers[first]->sync_shared_state(st[first].args, st[first].tx, st[first].rx); });
        std::this_thread::sleep_for(std::chrono::milliseconds(80));
        st[1 - first].ok = peers[1 - first]->sync_shared_state(st[1 - first].args, st[1 - first].tx, st[1 - first].rx);
        early.join();
        CHECK(st[0].ok && st[1].ok && st[0].buffers[0] == st[1].buffers[0]);
        if (first == 0) settled = st[0].buffers[0];
        else CHECK(settled == st[0].buffers[0]); // one vote each: the tie is broken by content, not by arrival
    }
}

// Several newcomers at once and several peers that hold the content: the master spreads the takers over the sources instead of
// sending everybody to the first one (the reference does the latter and says so in a comment, ccoip_master_handler.cpp:575).
TEST(takers_are_spread_over_the_sources) {
    TestMaster tm;
    auto peers = connect_peers(tm, 6);
    std::vector<PeerState> st(6);
    for (int p = 0; p < 6; ++p) st[p].add("w", 40'000, static_cast<uint8_t>(p < 3 ? 5 : 100 + p)); // three agree, three differ
    sync_all(peers, st);
    for (auto &s : st) CHECK(s.ok);
    for (int p = 3; p < 6; ++p) {
        CHECK_EQ(st[p].rx, 40'000u);
        CHECK(st[p].buffers[0] == st[0].buffers[0]);
    }
    for (int p = 0; p < 3; ++p) CHECK_EQ(st[p].tx, 40'000u); // one taker each
}

// SEND_ONLY towards two receivers: one distributor serves both, each receiver gets the full state exactly once
TEST(one_sender_two_receivers) {
    TestMaster tm;
    auto peers = connect_peers(tm, 3);
    std::vector<PeerState> st(3);
    st[0].add("a", 1000, 7);
    st[0].add("b", 33, 9, wire::DType::F32);
    for (int i = 1; i < 3; ++i) {
        st[i].add("a", 1000, static_cast<uint8_t>(20 + i));
        st[i].add("b", 33, static_cast<uint8_t>(30 + i), wire::DType::F32);
    }
    st[0].args.strategy = SyncStrategy::TxOnly;
    st[1].args.strategy = st[2].args.strategy = SyncStrategy::RxOnly;
    sync_all(peers, st);
    CHECK(st[0].ok && st[1].ok && st[2].ok);
    CHECK_EQ(st[0].tx, 2u * (1000 + 33 * 4));
    CHECK_EQ(st[1].rx, 1000u + 33 * 4);
    CHECK_EQ(st[2].rx, 1000u + 33 * 4);
    CHECK(st[1].buffers == st[0].buffers && st[2].buffers == st[0].buffers);
}

// a peer that joins between two syncs takes part from the next one
TEST(late_joiner_participates) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::vector<PeerState> st(3);
    for (auto &s : st) s.add("w", 777, 4, wire::DType::F16);
    st[0].args.revision = st[1].args.revision = 1;
    sync_all(peers, st, {0, 1});
    CHECK(st[0].ok && st[1].ok);
    // third peer connects while the others admit it
    auto late = std::make_unique<client::Client>(tm.config());
    std::atomic<bool> joined{false};
    std::thread joiner([&] {
        CHECK(late->connect());
        joined = true;
    });
    // keep admitting until the newcomer is in (once it is accepted it has to vote as well, so the
    // two old peers must stop voting on their own)
    while (peers[0]->global_world() < 3) {
        std::vector<std::thread> ts;
        for (auto &p : peers) ts.emplace_back([&p] { CHECK(p->update_topology()); });
        for (auto &t : ts) t.join();
    }
    joiner.join();
    CHECK(joined);
    peers.push_back(std::move(late));
    st[0].args.revision = st[1].args.revision = 2;
    st[2].args.revision = 2;
    std::fill(st[2].buffers[0].begin(), st[2].buffers[0].end(), uint8_t(99)); // newcomer holds garbage
    sync_all(peers, st);
    for (auto &s : st) CHECK(s.ok);
    CHECK_EQ(st[2].rx, 777u * 2);
    CHECK(st[2].buffers[0] == st[0].buffers[0]);
}

// Two leaders advance content and revision in lock step, a follower never touches its copy: every step the follower is
// outdated, receives the whole tensor from exactly one of the leaders and adopts their revision
// (reference test_shared_state_distribution.cpp:1318-1472)
TEST(follower_is_dragged_along_every_step) {
    TestMaster tm;
    auto peers = connect_peers(tm, 3);
    std::vector<PeerState> st(3);
    for (auto &s : st) s.add("w", 1500, 0, wire::DType::I16);
    for (int step = 0; step < 6; ++step) {
        for (int leader = 0; leader < 2; ++leader) {
            std::fill(st[leader].buffers[0].begin(), st[leader].buffers[0].end(), static_cast<uint8_t>(100 + step));
            st[leader].args.revision = 7 + step;
        }
        sync_all(peers, st);
        for (auto &s : st) CHECK(s.ok);
        CHECK_EQ(st[2].rx, 3000u);
        CHECK_EQ(st[2].tx, 0u);
        CHECK_EQ(st[0].rx + st[1].rx, 0u);
        CHECK((st[0].tx == 3000 && st[1].tx == 0) || (st[0].tx == 0 && st[1].tx == 3000));
        CHECK(st[2].buffers[0] == st[0].buffers[0]);
        CHECK_EQ(st[2].args.revision, uint64_t(7 + step));
    }
}

// The first revision a group ever sees may be anything (a run that resumes from a checkpoint); from then on the rule holds
// (reference :1582-1672)
TEST(first_revision_is_free_later_ones_are_not) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::vector<PeerState> st(2);
    for (auto &s : st) s.add("w", 64, 1);
    st[0].args.revision = st[1].args.revision = 13;
    sync_all(peers, st);
    CHECK(st[0].ok && st[1].ok);
    st[0].args.revision = st[1].args.revision = 14;
    sync_all(peers, st);
    CHECK(st[0].ok && st[1].ok);
    st[0].args.revision = st[1].args.revision = 16; // both skip one: both are kicked
    sync_all(peers, st);
    CHECK(!st[0].ok && !st[1].ok);
}

// Groups loop at their own pace on their own threads: ten steps each, one group twice as fast as the other; nothing moves,
// nobody waits for the other group (reference :1791-1934)
TEST(groups_advance_concurrently_at_their_own_pace) {
    TestMaster tm;
    auto peers = connect_peers(tm, 4, {0, 0, 1, 1});
    std::vector<PeerState> st(4);
    for (auto &s : st) s.add("w", 4096, 0);
    std::atomic<int> bad{0};
    auto run_group = [&](int a, int b, int delay_ms, uint8_t base) {
        for (int step = 0; step < 10; ++step) {
            for (int p : {a, b}) {
                std::fill(st[p].buffers[0].begin(), st[p].buffers[0].end(), static_cast<uint8_t>(base + step));
                st[p].args.revision = 1 + step;
            }
            std::thread other([&] { st[b].ok = peers[b]->sync_shared_state(st[b].args, st[b].tx, st[b].rx); });
            st[a].ok = peers[a]->sync_shared_state(st[a].args, st[a].tx, st[a].rx);
            other.join();
            for (int p : {a, b})
                if (!st[p].ok || st[p].tx || st[p].rx) bad++;
            std::this_thread::sleep_for(std::chrono::milliseconds(delay_ms));
        }
    };
    const auto t0 = std::chrono::steady_clock::now();
    std::chrono::steady_clock::duration fast_took{};
    std::thread g0([&] {
        run_group(0, 1, 5, 10);
        fast_took = std::chrono::steady_clock::now() - t0;
    });
    std::thread g1([&] { run_group(2, 3, 60, 200); });
    g0.join();
    g1.join();
    CHECK_EQ(bad.load(), 0);
    CHECK(st[0].buffers[0] == st[1].buffers[0] && st[2].buffers[0] == st[3].buffers[0]);
    CHECK(st[0].buffers[0] != st[2].buffers[0]);
    CHECK(fast_took < std::chrono::milliseconds(550)); // the slow group needs >= 600 ms for its sleeps alone
}

// Drag-along in three groups at once, each with two leaders and a follower and its own content (reference :1936-2080)
TEST(drag_along_in_several_groups_at_once) {
    TestMaster tm;
    auto peers = connect_peers(tm, 9, {0, 0, 0, 1, 1, 1, 2, 2, 2});
    std::vector<PeerState> st(9);
    for (auto &s : st) s.add("w", 2000, 0);
    std::atomic<int> bad{0};
    std::vector<std::thread> groups;
    for (int g = 0; g < 3; ++g)
        groups.emplace_back([&, g] {
            const int l0 = 3 * g, l1 = 3 * g + 1, f = 3 * g + 2;
            for (int step = 0; step < 4; ++step) {
                for (int p : {l0, l1}) {
                    std::fill(st[p].buffers[0].begin(), st[p].buffers[0].end(), static_cast<uint8_t>(50 * g + step + 1));
                    st[p].args.revision = 1 + step;
                }
                std::vector<std::thread> ts;
                for (int p : {l0, l1, f}) ts.emplace_back([&, p] { st[p].ok = peers[p]->sync_shared_state(st[p].args, st[p].tx, st[p].rx); });
                for (auto &t : ts) t.join();
                if (!(st[l0].ok && st[l1].ok && st[f].ok)) bad++;
                if (st[f].rx != 2000 || st[f].tx != 0 || st[l0].rx + st[l1].rx != 0 || st[l0].tx + st[l1].tx != 2000) bad++;
                if (st[f].buffers[0] != st[l0].buffers[0] || st[f].args.revision != uint64_t(1 + step)) bad++;
            }
        });
    for (auto &t : groups) t.join();
    CHECK_EQ(bad.load(), 0);
}

// A peer admitted between two synchronisation steps that already holds the right content and revision (it loaded the same
// checkpoint) takes part in the next step like everybody else: nothing moves (reference :2218-2378)
TEST(newcomer_between_steps_with_matching_state) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::vector<PeerState> st(3);
    for (auto &s : st) s.add("w", 999, 0);
    auto content = [&](int p, int step) {
        std::fill(st[p].buffers[0].begin(), st[p].buffers[0].end(), static_cast<uint8_t>(160 + step));
        st[p].args.revision = step;
    };
    for (int step = 0; step < 3; ++step) {
        content(0, step), content(1, step);
        sync_all(peers, st, {0, 1});
        CHECK(st[0].ok && st[1].ok && st[0].tx + st[1].tx + st[0].rx + st[1].rx == 0);
    }
    auto late = std::make_unique<client::Client>(tm.config());
    std::thread joiner([&] { CHECK(late->connect()); });
    while (peers[0]->global_world() < 3) {
        std::vector<std::thread> ts;
        for (auto &p : peers) ts.emplace_back([&p] { CHECK(p->update_topology()); });
        for (auto &t : ts) t.join();
    }
    joiner.join();
    peers.push_back(std::move(late));
    for (int step = 3; step < 6; ++step) {
        for (int p = 0; p < 3; ++p) content(p, step);
        sync_all(peers, st);
        for (auto &s : st) CHECK(s.ok && s.tx == 0 && s.rx == 0);
    }
}

// RECEIVE_ONLY everywhere with DIFFERENT content: still nobody to receive from, still everybody out (reference :2569-2664)
TEST(all_receive_only_with_different_content) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::vector<PeerState> st(2);
    st[0].add("w", 64, 1);
    st[1].add("w", 64, 2);
    st[0].args.strategy = st[1].args.strategy = SyncStrategy::RxOnly;
    sync_all(peers, st);
    CHECK(!st[0].ok && !st[1].ok);
    CHECK(st[0].buffers[0] == std::vector<uint8_t>(64, 1) && st[1].buffers[0] == std::vector<uint8_t>(64, 2)); // untouched
}

// A large state over sockets: every receiver fetches contiguous runs of the tensors over several connections at once, with the
// reference's plain request on each.  Sizes are uneven on purpose, one tensor is empty, one is larger than a run.
TEST(large_state_over_several_connections) {
    TestMaster tm;
    auto peers = connect_peers(tm, 3);
    std::vector<PeerState> st(3);
    const size_t sizes[] = {3u << 20, 1, 7u << 20, 0, 5u << 20, 30u << 20, 123457, 2u << 20, 9u << 20, 4u << 20, 64, 6u << 20};
    uint64_t total = 0;
    for (int p = 0; p < 3; ++p) {
        int k = 0;
        for (size_t n : sizes) st[p].add("tensor" + std::to_string(k++), n, static_cast<uint8_t>(p == 0 ? 0 : 200 + p));
    }
    for (size_t n : sizes) total += n;
    for (size_t k = 0; k < st[0].buffers.size(); ++k) // the sender's content: different in every tensor and along it
        for (size_t i = 0; i < st[0].buffers[k].size(); ++i) st[0].buffers[k][i] = static_cast<uint8_t>(i * 31 + k * 7 + (i >> 13));
    st[0].args.strategy = SyncStrategy::TxOnly;
    st[1].args.strategy = st[2].args.strategy = SyncStrategy::RxOnly;
    sync_all(peers, st);
    for (auto &s : st) CHECK(s.ok);
    CHECK_EQ(st[0].tx, 2 * total);
    CHECK_EQ(st[1].rx, total);
    CHECK_EQ(st[2].rx, total);
    for (int p = 1; p < 3; ++p)
        for (size_t k = 0; k < st[0].buffers.size(); ++k) CHECK(st[p].buffers[k] == st[0].buffers[k]);
    // the next synchronisation finds nothing to do
    for (auto &s : st) s.args.revision += 1, s.args.strategy = SyncStrategy::EnforcePopular;
    sync_all(peers, st);
    for (auto &s : st) CHECK(s.ok && s.tx == 0 && s.rx == 0);
}

// tensors with allow_content_inequality are never hashed or transferred
TEST(content_inequality_allowed) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::vector<PeerState> st(2);
    st[0].add("rng_state", 100, 1, wire::DType::U8, true);
    st[1].add("rng_state", 100, 2, wire::DType::U8, true);
    sync_all(peers, st);
    CHECK(st[0].ok && st[1].ok);
    CHECK_EQ(st[0].tx + st[1].tx, 0u);
    CHECK(st[0].buffers[0] != st[1].buffers[0]);
}

// First synchronisation of a group: the peer that resumes from a checkpoint (revision 5) wins over fresh peers
// (revision 0) no matter whose vote reaches the master first, and the fresh peers adopt its revision.
TEST(resumed_peer_wins_first_sync_in_any_vote_order) {
    for (int late : {0, 1, 2}) {
        TestMaster tm;
        auto peers = connect_peers(tm, 3);
        std::vector<PeerState> st(3);
        st[0].add("w", 4096, 42);
        st[0].args.revision = 5;
        st[1].add("w", 4096, 1);
        st[2].add("w", 4096, 1); // two fresh peers agree with each other: popularity alone would pick them
        std::vector<std::thread> ts;
        for (int i = 0; i < 3; ++i)
            ts.emplace_back([&, i] {
                if (i == late) std::this_thread::sleep_for(std::chrono::milliseconds(150));
                st[i].ok = peers[i]->sync_shared_state(st[i].args, st[i].tx, st[i].rx);
            });
        for (auto &t : ts) t.join();
        for (int i = 0; i < 3; ++i) {
            CHECK(st[i].ok);
            CHECK_EQ(st[i].buffers[0][0], 42);
            CHECK_EQ(st[i].buffers[0][4095], 42);
            CHECK_EQ(st[i].args.revision, 5u);
        }
        CHECK_EQ(st[0].rx, 0u);
        CHECK_EQ(st[1].rx, 4096u);
        CHECK_EQ(st[2].rx, 4096u);
    }
}

// popularity with several keys: the peer that votes first holds the minority content of ONE key and receives only that key
// (reference: test_shared_state_distribution.cpp:498)
TEST(popularity_is_decided_per_candidate_and_moves_only_dirty_keys) {
    TestMaster tm;
    auto peers = connect_peers(tm, 3);
    std::vector<PeerState> st(3);
    for (int p = 0; p < 3; ++p) {
        st[p].add("a", 300, 5);
        st[p].add("b", 700, p == 0 ? 99 : 6);
        st[p].add("c", 50, 7, wire::DType::F32);
    }
    std::thread first([&] { st[0].ok = peers[0]->sync_shared_state(st[0].args, st[0].tx, st[0].rx); });
    std::this_thread::sleep_for(std::chrono::milliseconds(100));
    sync_all(peers, st, {1, 2});
    first.join();
    CHECK(st[0].ok && st[1].ok && st[2].ok);
    CHECK_EQ(st[0].rx, 700u);
    CHECK(st[0].buffers[1] == std::vector<uint8_t>(700, 6));
    CHECK_EQ(st[1].rx + st[2].rx, 0u);
    CHECK_EQ(st[1].tx + st[2].tx, 700u);
}

// two groups synchronise at the same time, with the same key name meaning different things (size, content) in each:
// traffic stays inside the groups, nobody is kicked (reference: :771, :885, :2082)
TEST(groups_sync_concurrently_with_overlapping_key_names) {
    TestMaster tm;
    auto peers = connect_peers(tm, 4, {0, 0, 1, 1});
    std::vector<PeerState> st(4);
    st[0].add("w", 1000, 1);
    st[1].add("w", 1000, 2);
    st[2].add("w", 4000, 3, wire::DType::F32);
    st[3].add("w", 4000, 4, wire::DType::F32);
    for (int round = 0; round < 3; ++round) {
        sync_all(peers, st);
        for (auto &s : st) CHECK(s.ok);
        CHECK(st[0].buffers[0] == st[1].buffers[0]);
        CHECK(st[2].buffers[0] == st[3].buffers[0]);
        CHECK(st[0].buffers[0].size() == 1000 && st[2].buffers[0].size() == 16000);
        if (round == 0) {
            CHECK_EQ(st[0].rx + st[1].rx, 1000u);
            CHECK_EQ(st[2].rx + st[3].rx, 16000u);
        } else {
            for (auto &s : st) CHECK(s.tx == 0 && s.rx == 0);
        }
        for (auto &s : st) s.args.revision++;
    }
}

// the strategy-mixing rule is applied per group: {SEND_ONLY, SEND_ONLY} with equal content is fine in group 0 while
// group 1 = {ENFORCE_POPULAR, SEND_ONLY} loses its SEND_ONLY peer (reference: :3073); the kicked peer's later calls fail
TEST(strategy_mixing_rule_is_per_group) {
    TestMaster tm;
    auto peers = connect_peers(tm, 4, {0, 0, 1, 1});
    std::vector<PeerState> st(4);
    for (auto &s : st) s.add("w", 128, 9);
    st[0].args.strategy = SyncStrategy::TxOnly;
    st[1].args.strategy = SyncStrategy::TxOnly;
    st[2].args.strategy = SyncStrategy::EnforcePopular;
    st[3].args.strategy = SyncStrategy::TxOnly;
    sync_all(peers, st);
    CHECK(st[0].ok && st[1].ok && st[2].ok);
    CHECK(!st[3].ok);
    // the kicked peer is out for good
    uint64_t tx = 0, rx = 0;
    CHECK(!peers[3]->sync_shared_state(st[3].args, tx, rx));
    CHECK(!peers[3]->update_topology());
    // the others go on: group 0 keeps synchronising, the global world has shrunk
    for (int i = 0; i < 3; ++i) st[i].args.revision++;
    std::vector<std::thread> ts;
    for (int i = 0; i < 3; ++i) ts.emplace_back([&, i] { CHECK(peers[i]->update_topology()); });
    for (auto &t : ts) t.join();
    CHECK_EQ(peers[0]->global_world(), 3u);
    sync_all(peers, st, {0, 1});
    CHECK(st[0].ok && st[1].ok);
}

TEST_MAIN()
