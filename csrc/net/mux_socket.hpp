// One TCP connection carrying many logical streams, in ONE direction (a peer pair uses separate
// connections for each direction).  Frames are [u64 BE len = 16+n][u64 BE tag][u64 BE seq][n bytes];
// `seq` is the master-issued sequence number of the collective the bytes belong to, so a
// receiver can throw away left-overs of an aborted operation.  Behaviour model: reference
// tinysockets MultiplexedIOSocket (tinysockets/src/multiplexed_socket.cpp:253-714).
//
// Differences from the reference design: the TX queue entries reference caller memory
// (zero-copy; the caller waits on the ticket before touching the buffer), the receive side
// hands whole frame buffers to the consumer (which reduces straight out of them), and waiting
// consumers sleep on a condition variable instead of spinning.
#pragma once
#include <atomic>
#include <condition_variable>
#include <deque>
#include <memory>
#include <mutex>
#include <optional>
#include <thread>
#include <unordered_map>
#include <vector>

#include "../util/park.hpp"
#include "../util/ring_queue.hpp"
#include "../wire/packets.hpp"
#include "addr.hpp"

namespace pccl::net {

// Completion handle of one queued send: 0 in flight, 1 sent, 2 failed (reported as -1).
struct SendTicket {
    Parker word{0};
    void finish(bool ok) { word.publish(ok ? 1u : 2u); }
    // returns 1 sent, -1 failed, 0 on timeout
    int wait_for(int timeout_ms) {
        const uint32_t v = word.wait_while(0, timeout_ms);
        return v == 0 ? 0 : (v == 1 ? 1 : -1);
    }
};

class MuxSocket {
public:
    enum class Mode { Tx, Rx };
    MuxSocket(int fd, const ccoip_socket_address_t &remote, Mode mode);
    ~MuxSocket();
    MuxSocket(const MuxSocket &) = delete;
    MuxSocket &operator=(const MuxSocket &) = delete;

    bool run();
    bool interrupt();
    void join();
    bool is_open() const { return open_.load(std::memory_order_acquire); }
    const ccoip_socket_address_t &remote() const { return remote_; }

    // ---- TX side.  `data` must stay valid until the ticket finishes.  Large payloads are split
    // into frames of at most `frame_bytes`.
    std::shared_ptr<SendTicket> send_bytes(uint64_t tag, uint64_t seq, const void *data, size_t n, size_t frame_bytes);
    // Small owned payload (P2P packets: [u16 id][body]).
    template <typename P> bool send_packet(uint64_t tag, uint64_t seq, const P &pkt) {
        wire::Writer w;
        w.put<uint16_t>(P::id);
        pkt.write(w);
        return send_owned(tag, seq, std::move(w.bytes()));
    }
    bool send_owned(uint64_t tag, uint64_t seq, std::vector<uint8_t> bytes);

    // A received frame.  The buffer is deliberately NOT value-initialised (a std::vector would
    // zero every byte before the socket overwrites it).
    struct Frame {
        std::unique_ptr<uint8_t[]> bytes;
        size_t n = 0;
        const uint8_t *data() const { return bytes.get(); }
        size_t size() const { return n; }
    };
    // ---- RX side.  Pops the oldest frame of `tag` whose sequence number equals `seq`; frames
    // with another sequence number are dropped on the way.  timeout_ms == 0 polls.
    std::optional<Frame> recv_frame(uint64_t tag, uint64_t seq, int timeout_ms);
    template <typename P> std::optional<P> recv_packet(uint64_t tag, uint64_t seq, int timeout_ms) {
        auto frame = recv_frame(tag, seq, timeout_ms);
        if (!frame || frame->size() < 2) return std::nullopt;
        wire::Reader r(frame->data(), frame->size());
        if (r.get<uint16_t>() != P::id) return std::nullopt;
        P pkt;
        if (!pkt.read(r)) return std::nullopt;
        return pkt;
    }
    void purge_tag(uint64_t tag);

private:
    struct TxItem {
        uint64_t tag, seq;
        const uint8_t *data;
        size_t n, frame_bytes;
        std::vector<uint8_t> owned;
        std::shared_ptr<SendTicket> ticket;
    };
    struct RxFrame {
        uint64_t seq;
        Frame data;
    };
    void tx_loop();
    void rx_loop();
    void fail_all_tx();

    int fd_;
    ccoip_socket_address_t remote_;
    Mode mode_;
    std::atomic<bool> open_{true}, stop_{false};
    std::thread thread_;

    // send requests: lock-free ring filled by any thread, drained by the TX thread, which sleeps on `tx_signal_`
    static constexpr size_t kTxDepth = 1024;
    MpscRing<TxItem> tx_queue_{kTxDepth};
    Parker tx_signal_{0};
    std::atomic<bool> tx_exited_{false};
    std::mutex tx_drain_mu_;
    bool enqueue(TxItem &&item);

    std::mutex rx_mu_;
    std::condition_variable rx_cv_;
    std::unordered_map<uint64_t, std::deque<RxFrame>> rx_streams_;
};

} // namespace pccl::net
