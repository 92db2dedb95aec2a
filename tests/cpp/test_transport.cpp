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
