// Transport behaviours (reference tinysockets/tests/test_server_socket.cpp,
// test_blocking_io_socket.cpp, test_queued_socket.cpp) over real loopback sockets.
#include <atomic>
#include <chrono>
#include <condition_variable>
#include <map>
#include <mutex>
#include <thread>

#include "harness.hpp"
#include "net/blocking_socket.hpp"
#include "net/epoll_server.hpp"
#include "net/mux_socket.hpp"
#include "net/queued_socket.hpp"
#include "net/sockutil.hpp"

using namespace pccl;
using namespace pccl::net;
using namespace std::chrono_literals;

static ccoip_socket_address_t loopback(uint16_t port) {
    ccoip_socket_address_t a{};
    a.inet.protocol = inetIPv4;
    a.inet.ipv4 = {{127, 0, 0, 1}};
    a.port = port;
    return a;
}

template <typename F> static bool eventually(F f, int ms = 3000) {
    for (int i = 0; i < ms; ++i) {
        if (f()) return true;
        std::this_thread::sleep_for(1ms);
    }
    return false;
}

// raw frame with an arbitrary packet id
struct RawPacket {
    static constexpr wire::PacketId id = 999;
    std::vector<uint8_t> body;
    void write(wire::Writer &w) const { w.put_bytes(body.data(), body.size()); }
    bool read(wire::Reader &r) {
        body.resize(r.remaining());
        return r.get_bytes(body.data(), body.size());
    }
};

struct OddPacket { // an id no packet of the protocol has
    static constexpr wire::PacketId id = 0x7abc;
    void write(wire::Writer &w) const { w.put<uint32_t>(0xdeadbeef); }
};

TEST(server_lifecycle) {
    EpollServer s(loopback(0));
    CHECK(!s.run_async()); // not listening yet
    CHECK(s.listen());
    CHECK(!s.listen()); // twice fails
    CHECK(s.port() != 0);
    CHECK(s.run_async());
    CHECK(!s.run_async());
    CHECK(s.interrupt());
    CHECK(s.interrupt()); // idempotent
    s.join();
    // a second server cannot take a busy port, but a bumping one moves on
    EpollServer a(loopback(0));
    CHECK(a.listen());
    EpollServer clash(loopback(a.port()));
    CHECK(!clash.listen());
    EpollServer bump(loopback(a.port()), true);
    CHECK(bump.listen());
    CHECK(bump.port() != a.port());
}

TEST(server_frames_and_callbacks) {
    EpollServer s(loopback(0));
    std::mutex mu;
    std::vector<std::pair<uint16_t, size_t>> seen;
    std::atomic<int> joins{0}, closes{0};
    s.on_join([&](const ccoip_socket_address_t &) { joins++; });
    s.on_close([&](const ccoip_socket_address_t &) { closes++; });
    s.on_read([&](const ccoip_socket_address_t &from, std::span<const uint8_t> frame) {
        wire::Reader r(frame.data(), frame.size());
        const uint16_t id = r.get<uint16_t>();
        {
            std::lock_guard<std::mutex> lock(mu);
            seen.emplace_back(id, r.remaining());
        }
        if (id == 999) { // echo back
            RawPacket p;
            p.read(r);
            s.send_packet(from, p);
        }
        if (id == 7) s.close_client(from); // kick
    });
    CHECK(s.listen() && s.run_async());
    BlockingSocket c(loopback(s.port()));
    CHECK(c.connect());
    CHECK(eventually([&] { return joins == 1; }));
    // zero-length payloads are legal
    CHECK(c.send_packet(wire::C2MCheckPeersPending{}));
    // a 10 MB frame is reassembled and echoed
    RawPacket big;
    big.body.resize(10u << 20);
    for (size_t i = 0; i < big.body.size(); ++i) big.body[i] = static_cast<uint8_t>(i * 31);
    CHECK(c.send_packet(big));
    const auto echo = c.recv_packet<RawPacket>(10000);
    CHECK(echo && echo->body == big.body);
    CHECK(eventually([&] {
        std::lock_guard<std::mutex> lock(mu);
        return seen.size() == 2;
    }));
    CHECK(seen[0].first == wire::C2MCheckPeersPending::id && seen[0].second == 0);
    // frames split into tiny writes still arrive whole
    const auto frame = wire::encode_frame(big);
    for (size_t off = 0; off < 64; ++off) CHECK(c.send_raw(frame.data() + off, 1));
    CHECK(c.send_raw(frame.data() + 64, frame.size() - 64));
    CHECK(c.recv_packet<RawPacket>(10000).has_value());
    // kicked client: exactly one close callback, connection really closed
    wire::C2MOptimizeWorkComplete kick_me; // id 7
    CHECK(c.send_packet(kick_me));
    CHECK(eventually([&] { return closes == 1; }));
    CHECK(eventually([&] { return !c.is_open(); }));
    std::this_thread::sleep_for(50ms);
    CHECK_EQ(closes.load(), 1);
    // an oversized length prefix gets the sender dropped
    BlockingSocket evil(loopback(s.port()));
    CHECK(evil.connect());
    const uint8_t huge[8] = {0x7f, 0, 0, 0, 0, 0, 0, 0};
    CHECK(evil.send_raw(huge, 8));
    CHECK(eventually([&] { return closes == 2; }));
    s.interrupt();
    s.join();
}

TEST(server_many_concurrent_clients) {
    EpollServer s(loopback(0));
    std::atomic<uint64_t> bytes{0};
    s.on_read([&](const ccoip_socket_address_t &from, std::span<const uint8_t> frame) {
        bytes += frame.size() - 2;
        RawPacket ack;
        ack.body = {1};
        s.send_packet(from, ack);
    });
    CHECK(s.listen() && s.run_async());
    constexpr int kClients = 16, kFrames = 20;
    std::vector<std::thread> ts;
    for (int i = 0; i < kClients; ++i)
        ts.emplace_back([&, i] {
            BlockingSocket c(loopback(s.port()));
            CHECK(c.connect());
            RawPacket p;
            p.body.assign(100000 + i, static_cast<uint8_t>(i));
            for (int k = 0; k < kFrames; ++k) {
                CHECK(c.send_packet(p));
                CHECK(c.recv_packet<RawPacket>(10000).has_value());
            }
        });
    for (auto &t : ts) t.join();
    uint64_t want = 0;
    for (int i = 0; i < kClients; ++i) want += uint64_t(kFrames) * (100000 + i);
    CHECK_EQ(bytes.load(), want);
    s.interrupt();
    s.join();
}

// many clients, each with several multi-megabyte frames in flight at once, interleaved on the wire: every frame is
// reassembled intact and in per-connection order (reference tinysockets/tests/test_server_socket.cpp:1074-1186)
TEST(server_many_concurrent_large_frames) {
    EpollServer s(loopback(0));
    std::mutex mu;
    std::map<uint8_t, std::vector<uint32_t>> seen; // client id -> frame numbers in arrival order
    std::atomic<int> bad{0};
    s.on_read([&](const ccoip_socket_address_t &, std::span<const uint8_t> frame) {
        // frame = [u16 id][client][frame number u32 LE][payload of a size that depends on both]
        if (frame.size() < 7) {
            ++bad;
            return;
        }
        const uint8_t client = frame[2];
        uint32_t number = 0;
        std::memcpy(&number, frame.data() + 3, 4);
        const size_t want = 7 + (3u << 20) + 4099u * client + 17u * number;
        if (frame.size() != want) ++bad;
        for (size_t i = 7; i < frame.size(); i += 997)
            if (frame[i] != static_cast<uint8_t>(client * 31 + number + i)) ++bad;
        std::lock_guard<std::mutex> lock(mu);
        seen[client].push_back(number);
    });
    CHECK(s.listen() && s.run_async());
    constexpr int kClients = 8, kFrames = 6;
    std::vector<std::thread> ts;
    for (int c = 0; c < kClients; ++c)
        ts.emplace_back([&, c] {
            BlockingSocket sock(loopback(s.port()));
            CHECK(sock.connect());
            for (uint32_t k = 0; k < kFrames; ++k) {
                RawPacket p;
                p.body.resize(5 + (3u << 20) + 4099u * c + 17u * k);
                p.body[0] = static_cast<uint8_t>(c);
                std::memcpy(p.body.data() + 1, &k, 4);
                for (size_t i = 7; i < p.body.size() + 2; ++i) p.body[i - 2] = static_cast<uint8_t>(c * 31 + k + i);
                CHECK(sock.send_packet(p)); // no acknowledgement in between: frames pile up behind each other
            }
        });
    for (auto &t : ts) t.join();
    CHECK(eventually([&] {
        std::lock_guard<std::mutex> lock(mu);
        size_t n = 0;
        for (auto &[c, v] : seen) n += v.size();
        return n == size_t(kClients) * kFrames;
    }));
    CHECK_EQ(bad.load(), 0);
    for (auto &[c, v] : seen)
        for (uint32_t k = 0; k < kFrames; ++k) CHECK_EQ(v[k], k);
    s.interrupt();
    s.join();
}

// addresses a server can and cannot take (reference test_server_socket.cpp: test_bind_invalid, test_bind_valid_ipv6,
// test_bind_invalid_ipv6, test_client_connect_when_not_listening)
TEST(server_addresses) {
    ccoip_socket_address_t not_ours{};
    not_ours.inet.protocol = inetIPv4;
    not_ours.inet.ipv4 = {{203, 0, 113, 7}}; // TEST-NET-3: no interface of this machine has it
    EpollServer bad(not_ours);
    CHECK(!bad.listen());
    CHECK(!bad.run_async());
    CHECK(bad.listen() == false); // a failed listen can be tried again and fails again, it does not wedge the object
    ccoip_socket_address_t v6{};
    v6.inet.protocol = inetIPv6;
    v6.inet.ipv6.data[15] = 1; // ::1
    EpollServer six(v6);
    if (six.listen()) { // the sandbox may have no IPv6 loopback
        CHECK(six.port() != 0);
        CHECK(six.run_async());
        ccoip_socket_address_t to = v6;
        to.port = six.port();
        BlockingSocket c(to);
        CHECK(c.connect());
        six.interrupt();
        six.join();
    }
    ccoip_socket_address_t bad6{};
    bad6.inet.protocol = inetIPv6;
    bad6.inet.ipv6.data[0] = 0x20, bad6.inet.ipv6.data[1] = 0x01, bad6.inet.ipv6.data[2] = 0x0d, bad6.inet.ipv6.data[3] = 0xb8; // 2001:db8::, documentation prefix
    EpollServer nowhere6(bad6);
    CHECK(!nowhere6.listen());
    // bound but not yet running: the kernel completes the handshake (backlog), nothing is read until the loop runs
    EpollServer idle(loopback(0));
    BlockingSocket early(loopback(1));
    CHECK(!early.connect(300)); // nobody there at all
}

// what stopping means (reference: test_run_async_after_interrupt, test_rerun_after_join, test_multiple_interrupts,
// test_client_send_after_server_stop, test_close_client_after_server_stopped, test_close_non_existent_client,
// test_server_send_to_non_existent_clients)
TEST(server_stop_semantics) {
    EpollServer s(loopback(0));
    ccoip_socket_address_t client{};
    std::atomic<bool> have{false};
    s.on_join([&](const ccoip_socket_address_t &a) {
        client = a;
        have = true;
    });
    CHECK(s.listen() && s.run_async());
    ccoip_socket_address_t ghost = loopback(9);
    CHECK(!s.close_client(ghost));
    RawPacket p;
    p.body = {1, 2, 3};
    CHECK(!s.send_packet(ghost, p));
    BlockingSocket c(loopback(s.port()));
    CHECK(c.connect());
    CHECK(eventually([&] { return have.load(); }));
    CHECK(s.interrupt());
    CHECK(s.interrupt());
    s.join();
    s.join(); // joining twice is harmless
    CHECK(!s.run_async()); // a stopped server stays stopped
    CHECK(!s.listen());
    // the stopped server closed its connections: the client finds out at the latest when it reads
    CHECK(eventually([&] {
        c.send_packet(p);
        return !c.recv_packet<RawPacket>(20).has_value() && !c.is_open();
    }));
    CHECK(!s.close_client(client));
    CHECK(!s.send_packet(client, p));
}

// the server talks first, talks back, and notices that a client has gone (reference: test_server_send_on_client_connect,
// test_server_send_on_client_message, test_server_send_after_client_disconnect, test_server_send_large_packets,
// test_server_send_to_multiple_clients)
TEST(server_sends) {
    EpollServer s(loopback(0));
    std::mutex mu;
    std::vector<ccoip_socket_address_t> clients;
    std::atomic<int> closes{0};
    s.on_join([&](const ccoip_socket_address_t &a) {
        RawPacket hello;
        hello.body = {'h', 'i'};
        CHECK(s.send_packet(a, hello)); // from inside the join callback
        std::lock_guard<std::mutex> lock(mu);
        clients.push_back(a);
    });
    s.on_close([&](const ccoip_socket_address_t &) { closes++; });
    s.on_read([&](const ccoip_socket_address_t &from, std::span<const uint8_t> frame) {
        RawPacket twice; // two answers per message, from inside the read callback
        twice.body.assign(frame.begin() + 2, frame.end());
        CHECK(s.send_packet(from, twice));
        twice.body.push_back(0xff);
        CHECK(s.send_packet(from, twice));
    });
    CHECK(s.listen() && s.run_async());
    constexpr int kClients = 5;
    std::vector<std::unique_ptr<BlockingSocket>> socks;
    for (int i = 0; i < kClients; ++i) {
        socks.push_back(std::make_unique<BlockingSocket>(loopback(s.port())));
        CHECK(socks.back()->connect());
        const auto hi = socks.back()->recv_packet<RawPacket>(3000);
        CHECK(hi && hi->body.size() == 2);
    }
    for (int i = 0; i < kClients; ++i) {
        RawPacket q;
        q.body.assign(1000 + i, static_cast<uint8_t>(i));
        CHECK(socks[i]->send_packet(q));
    }
    for (int i = 0; i < kClients; ++i) {
        const auto a = socks[i]->recv_packet<RawPacket>(3000), b = socks[i]->recv_packet<RawPacket>(3000);
        CHECK(a && b && a->body.size() == size_t(1000 + i) && b->body.size() == size_t(1001 + i) && b->body.back() == 0xff);
    }
    // unsolicited large packets from another thread, to everyone at once, while the clients read slowly
    RawPacket big;
    big.body.resize(9u << 20);
    for (size_t i = 0; i < big.body.size(); ++i) big.body[i] = static_cast<uint8_t>(i ^ (i >> 11));
    {
        std::lock_guard<std::mutex> lock(mu);
        CHECK_EQ(clients.size(), size_t(kClients));
        for (int round = 0; round < 2; ++round)
            for (const auto &a : clients) CHECK(s.send_packet(a, big));
    }
    std::vector<std::thread> readers;
    std::atomic<int> good{0};
    for (int i = 0; i < kClients; ++i)
        readers.emplace_back([&, i] {
            for (int round = 0; round < 2; ++round) {
                std::this_thread::sleep_for(std::chrono::milliseconds(5 * i));
                const auto got = socks[i]->recv_packet<RawPacket>(20000);
                if (got && got->body == big.body) good++;
            }
        });
    for (auto &t : readers) t.join();
    CHECK_EQ(good.load(), 2 * kClients);
    // a client that left cannot be written to any more, the others can
    ccoip_socket_address_t gone;
    {
        std::lock_guard<std::mutex> lock(mu);
        gone = clients[0];
    }
    socks[0]->close();
    CHECK(eventually([&] { return closes == 1; }));
    RawPacket small;
    small.body = {7};
    CHECK(!s.send_packet(gone, small));
    {
        std::lock_guard<std::mutex> lock(mu);
        CHECK(s.send_packet(clients[1], small));
    }
    CHECK(socks[1]->recv_packet<RawPacket>(3000).has_value());
    s.interrupt();
    s.join();
}

// edge sizes (reference: test_client_send_exact_max_packet_size, test_client_send_extremely_large_packet,
// test_unknown_packet_id): the largest frame the server accepts arrives whole, one byte more costs the connection, and an
// id nobody knows is handed to the application untouched -- deciding what it means is not the transport's business.
TEST(server_frame_size_limits_and_unknown_ids) {
    EpollServer s(loopback(0));
    std::atomic<uint64_t> last_size{0};
    std::atomic<int> frames{0}, closes{0};
    std::atomic<uint16_t> last_id{0};
    s.on_read([&](const ccoip_socket_address_t &, std::span<const uint8_t> frame) {
        wire::Reader r(frame.data(), frame.size());
        last_id = r.get<uint16_t>();
        last_size = frame.size();
        frames++;
    });
    s.on_close([&](const ccoip_socket_address_t &) { closes++; });
    CHECK(s.listen() && s.run_async());
    BlockingSocket c(loopback(s.port()));
    CHECK(c.connect());
    RawPacket exact;
    exact.body.assign(wire::kMaxControlFrame - 2, 0x5a); // + the two id bytes = the cap exactly
    CHECK(c.send_packet(exact));
    CHECK(eventually([&] { return frames == 1; }, 20000));
    CHECK_EQ(last_size.load(), wire::kMaxControlFrame);
    CHECK_EQ(last_id.load(), uint16_t(999));
    CHECK(c.send_packet(OddPacket{}));
    CHECK(eventually([&] { return frames == 2; }));
    CHECK_EQ(last_id.load(), uint16_t(0x7abc));
    CHECK_EQ(closes.load(), 0);
    RawPacket over;
    over.body.assign(wire::kMaxControlFrame - 1, 0);
    c.send_packet(over); // may fail half way: the server hangs up as soon as it has seen the length
    CHECK(eventually([&] { return closes == 1; }, 20000));
    CHECK_EQ(frames.load(), 2);
    s.interrupt();
    s.join();
}

// predicate-matched receive never loses or reorders packets that do not match
TEST(queued_socket_matching) {
    EpollServer s(loopback(0));
    ccoip_socket_address_t client{};
    std::atomic<bool> have_client{false};
    s.on_join([&](const ccoip_socket_address_t &a) {
        client = a;
        have_client = true;
    });
    CHECK(s.listen() && s.run_async());
    QueuedSocket q(loopback(s.port()));
    CHECK(q.connect() && q.run());
    CHECK(eventually([&] { return have_client.load(); }));
    // nothing there yet
    CHECK(!q.recv_packet<wire::M2CCollectiveCommence>(true).has_value());
    for (uint64_t tag = 0; tag < 50; ++tag) {
        wire::M2CCollectiveCommence c;
        c.tag = tag, c.seq_nr = 1000 + tag;
        CHECK(s.send_packet(client, c));
        wire::M2CCollectiveAbort a;
        a.tag = tag, a.aborted = tag % 2;
        CHECK(s.send_packet(client, a));
    }
    // consumers on many threads, each waiting for its own tag, in reverse order
    std::vector<std::thread> ts;
    std::atomic<int> ok{0};
    for (uint64_t tag = 0; tag < 50; ++tag)
        ts.emplace_back([&, tag] {
            const auto c = q.recv_matching<wire::M2CCollectiveCommence>([&](const wire::M2CCollectiveCommence &p) { return p.tag == 49 - tag; });
            if (c && c->seq_nr == 1000 + 49 - tag) ok++;
        });
    for (auto &t : ts) t.join();
    CHECK_EQ(ok.load(), 50);
    // the abort packets are all still there, in order
    for (uint64_t tag = 0; tag < 50; ++tag) {
        const auto a = q.recv_packet<wire::M2CCollectiveAbort>();
        CHECK(a && a->tag == tag && a->aborted == (tag % 2 == 1));
    }
    // a blocked receiver is released when the link dies
    std::thread waiter([&] { CHECK(!q.recv_packet<wire::M2CCollectiveComplete>().has_value()); });
    std::this_thread::sleep_for(50ms);
    s.close_client(client);
    waiter.join();
    CHECK(!q.is_open());
    CHECK(!q.send_packet(wire::C2MCheckPeersPending{}));
    s.interrupt();
    s.join();
}

// producers and consumers at once: many threads wait for their own packet while the packets are still arriving, interleaved
// with packets of another type; nothing is lost, nothing is delivered twice, and the observer saw every frame in wire order
// (reference test_queued_socket.cpp: *_multiple_threads__many_packets, test_interleaved_packet_types)
TEST(queued_socket_many_threads_many_packets) {
    EpollServer s(loopback(0));
    ccoip_socket_address_t client{};
    std::atomic<bool> have_client{false};
    s.on_join([&](const ccoip_socket_address_t &a) {
        client = a;
        have_client = true;
    });
    CHECK(s.listen() && s.run_async());
    QueuedSocket q(loopback(s.port()));
    std::vector<uint16_t> wire_order;
    q.set_observer([&](uint16_t id, const std::vector<uint8_t> &) { wire_order.push_back(id); }); // receive thread only
    CHECK(q.connect() && q.run());
    CHECK(eventually([&] { return have_client.load(); }));
    constexpr uint64_t kTags = 16, kPerTag = 64;
    std::vector<std::thread> consumers;
    std::atomic<uint64_t> commences{0}, completes{0}, wrong{0};
    for (uint64_t tag = 0; tag < kTags; ++tag)
        consumers.emplace_back([&, tag] {
            uint64_t last = 0;
            for (uint64_t k = 0; k < kPerTag; ++k) {
                const auto c = q.recv_matching<wire::M2CCollectiveCommence>([&](const wire::M2CCollectiveCommence &p) { return p.tag == tag; });
                if (!c || (k > 0 && c->seq_nr <= last)) wrong++; // per tag, arrival order is delivery order
                last = c ? c->seq_nr : 0;
                commences++;
            }
        });
    std::thread other([&] {
        for (uint64_t k = 0; k < kTags * kPerTag; ++k) {
            const auto d = q.recv_packet<wire::M2CCollectiveComplete>();
            if (!d) wrong++;
            completes++;
        }
    });
    std::thread producer([&] {
        for (uint64_t k = 0; k < kPerTag; ++k)
            for (uint64_t tag = 0; tag < kTags; ++tag) {
                wire::M2CCollectiveCommence c;
                c.tag = kTags - 1 - tag, c.seq_nr = 1 + k;
                CHECK(s.send_packet(client, c));
                wire::M2CCollectiveComplete d;
                d.tag = tag;
                CHECK(s.send_packet(client, d));
            }
    });
    producer.join();
    for (auto &t : consumers) t.join();
    other.join();
    CHECK_EQ(commences.load(), kTags * kPerTag);
    CHECK_EQ(completes.load(), kTags * kPerTag);
    CHECK_EQ(wrong.load(), uint64_t(0));
    CHECK(!q.recv_packet<wire::M2CCollectiveCommence>(true).has_value());
    CHECK(!q.recv_packet<wire::M2CCollectiveComplete>(true).has_value());
    q.interrupt();
    q.join();
    CHECK_EQ(wire_order.size(), size_t(2 * kTags * kPerTag));
    for (size_t i = 0; i < wire_order.size(); ++i)
        CHECK_EQ(wire_order[i], i % 2 == 0 ? wire::M2CCollectiveCommence::id : wire::M2CCollectiveComplete::id);
    s.interrupt();
    s.join();
}

// Two blocking sockets talking to each other in both directions at once with frames far larger than the socket buffers
// (reference test_blocking_io_socket.cpp: test_send_and_receive_large_packets_concurrently, test_concurrent_clients_send_large_packets)
TEST(blocking_sockets_full_duplex_large_frames) {
    Listener l(loopback(0), false);
    std::mutex mu;
    std::condition_variable cv;
    std::vector<std::unique_ptr<BlockingSocket>> accepted;
    l.on_join([&](std::unique_ptr<BlockingSocket> s) {
        std::lock_guard<std::mutex> lock(mu);
        accepted.push_back(std::move(s));
        cv.notify_all();
    });
    CHECK(l.listen() && l.run_async());
    constexpr int kPairs = 4, kFrames = 3;
    std::vector<std::unique_ptr<BlockingSocket>> dialled;
    for (int i = 0; i < kPairs; ++i) {
        dialled.push_back(std::make_unique<BlockingSocket>(loopback(l.port())));
        CHECK(dialled.back()->connect());
        // tell the other end which pair this is: accept order is not connect order
        RawPacket who;
        who.body = {static_cast<uint8_t>(i)};
        CHECK(dialled.back()->send_packet(who));
    }
    {
        std::unique_lock<std::mutex> lock(mu);
        CHECK(cv.wait_for(lock, 5s, [&] { return accepted.size() == size_t(kPairs); }));
    }
    auto payload = [](int pair, int dir, int k) {
        RawPacket p;
        p.body.resize((12u << 20) + 4097u * pair + 13u * k + dir);
        for (size_t i = 0; i < p.body.size(); i += 61) p.body[i] = static_cast<uint8_t>(pair * 7 + dir * 3 + k + i);
        return p;
    };
    std::atomic<int> good{0};
    std::vector<std::thread> ts;
    for (int i = 0; i < kPairs; ++i) {
        ts.emplace_back([&, i] { // dialling side: writes, and reads on a second thread
            std::thread reader([&] {
                for (int k = 0; k < kFrames; ++k) {
                    const auto got = dialled[i]->recv_packet<RawPacket>(30000);
                    if (got && got->body == payload(i, 1, k).body) good++;
                }
            });
            for (int k = 0; k < kFrames; ++k) CHECK(dialled[i]->send_packet(payload(i, 0, k)));
            reader.join();
        });
        ts.emplace_back([&, i] { // accepting side
            BlockingSocket &sock = *accepted[i];
            const auto who = sock.recv_packet<RawPacket>(5000);
            CHECK(who && who->body.size() == 1);
            const int pair = who ? who->body[0] : 0;
            std::thread reader([&] {
                for (int k = 0; k < kFrames; ++k) {
                    const auto got = sock.recv_packet<RawPacket>(30000);
                    if (got && got->body == payload(pair, 0, k).body) good++;
                }
            });
            for (int k = 0; k < kFrames; ++k) CHECK(sock.send_packet(payload(pair, 1, k)));
            reader.join();
        });
    }
    for (auto &t : ts) t.join();
    CHECK_EQ(good.load(), 2 * kPairs * kFrames);
    l.interrupt();
    l.join();
}

TEST(multiplexed_streams_and_stale_sequence_numbers) {
    Listener l(loopback(0), false);
    std::shared_ptr<MuxSocket> rx;
    std::mutex mu;
    std::condition_variable cv;
    l.on_join([&](std::unique_ptr<BlockingSocket> s) {
        auto m = std::make_shared<MuxSocket>(s->release(), s->remote(), MuxSocket::Mode::Rx);
        m->run();
        std::lock_guard<std::mutex> lock(mu);
        rx = m;
        cv.notify_all();
    });
    CHECK(l.listen() && l.run_async());
    const int fd = connect_tcp(loopback(l.port()));
    CHECK(fd >= 0);
    MuxSocket tx(fd, loopback(l.port()), MuxSocket::Mode::Tx);
    CHECK(tx.run());
    {
        std::unique_lock<std::mutex> lock(mu);
        cv.wait_for(lock, 3s, [&] { return rx != nullptr; });
    }
    CHECK(rx != nullptr);
    std::vector<uint8_t> a(3u << 20), b(1000), stale(500, 0xee);
    for (size_t i = 0; i < a.size(); ++i) a[i] = static_cast<uint8_t>(i * 7);
    for (size_t i = 0; i < b.size(); ++i) b[i] = static_cast<uint8_t>(i);
    // leftovers of an aborted operation (old sequence number) precede the real data on tag 1
    auto t0 = tx.send_bytes(1, 41, stale.data(), stale.size(), 256);
    auto t1 = tx.send_bytes(1, 42, a.data(), a.size(), 1u << 20); // split into 3 frames
    auto t2 = tx.send_bytes(2, 42, b.data(), b.size(), 0);
    auto t3 = tx.send_bytes(3, 42, nullptr, 0, 0); // zero-length payloads travel too
    CHECK(t0->wait_for(5000) == 1 && t1->wait_for(5000) == 1 && t2->wait_for(5000) == 1 && t3->wait_for(5000) == 1);
    // tag 2 can be read before tag 1
    const auto fb = rx->recv_frame(2, 42, 3000);
    CHECK(fb && fb->size() == b.size() && std::memcmp(fb->data(), b.data(), b.size()) == 0);
    std::vector<uint8_t> got;
    while (got.size() < a.size()) {
        const auto f = rx->recv_frame(1, 42, 3000);
        CHECK(f.has_value());
        CHECK(f->size() <= (1u << 20));
        got.insert(got.end(), f->data(), f->data() + f->size());
    }
    CHECK(got == a); // the stale frames were discarded, order preserved
    const auto fz = rx->recv_frame(3, 42, 3000);
    CHECK(fz && fz->size() == 0);
    CHECK(!rx->recv_frame(1, 42, 0).has_value());
    // P2P packets ride on the same streams
    wire::P2PHello hello;
    hello.uuid = Uuid::random_v4();
    hello.connection_nr = 3;
    CHECK(tx.send_packet(9, 1, hello));
    const auto h = rx->recv_packet<wire::P2PHello>(9, 1, 3000);
    CHECK(h && h->uuid == hello.uuid && h->connection_nr == 3);
    // closing the sender ends the receiver; later sends fail immediately
    tx.interrupt();
    CHECK(eventually([&] { return !rx->is_open(); }));
    CHECK(tx.send_bytes(1, 43, b.data(), b.size(), 0)->wait_for(100) == -1);
    l.interrupt();
    l.join();
}

TEST(blocking_socket_basics) {
    BlockingSocket nowhere(loopback(1)); // nothing listens on port 1
    CHECK(!nowhere.connect(500));
    Listener l(loopback(0), false);
    std::atomic<int> accepted{0};
    l.on_join([&](std::unique_ptr<BlockingSocket> s) {
        accepted++;
        const auto p = s->recv_packet<wire::P2PHello>(2000);
        if (p) s->send_packet(wire::P2PHelloAck{});
        // a packet of the wrong type is reported as absent
        s->recv_packet<wire::P2PHello>(200);
    });
    CHECK(l.listen() && l.run_async());
    BlockingSocket c(loopback(l.port()));
    CHECK(c.connect());
    CHECK(c.is_open());
    CHECK(!c.recv_packet<wire::P2PHelloAck>(50).has_value()); // timeout, nothing sent yet
    wire::P2PHello hello;
    CHECK(c.send_packet(hello));
    CHECK(c.recv_packet<wire::P2PHelloAck>(2000).has_value());
    CHECK(eventually([&] { return accepted == 1; }));
    c.close();
    CHECK(!c.is_open());
    l.interrupt();
    l.join();
}

TEST_MAIN()
