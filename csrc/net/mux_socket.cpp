#include "mux_socket.hpp"

#include <sys/uio.h>
#include <unistd.h>

#include <cerrno>

#include "../util/log.hpp"
#include "sockutil.hpp"

namespace pccl::net {

MuxSocket::MuxSocket(int fd, const ccoip_socket_address_t &remote, Mode mode) : fd_(fd), remote_(remote), mode_(mode) {
    set_keepalive(fd_, 30);
}

MuxSocket::~MuxSocket() {
    interrupt();
    join();
    if (fd_ >= 0) ::close(fd_);
}

bool MuxSocket::run() {
    if (thread_.joinable()) return false;
    thread_ = std::thread([this] { mode_ == Mode::Tx ? tx_loop() : rx_loop(); });
    return true;
}

bool MuxSocket::interrupt() {
    stop_ = true;
    open_.store(false, std::memory_order_release);
    if (fd_ >= 0) ::shutdown(fd_, SHUT_RDWR);
    tx_signal_.bump();
    rx_cv_.notify_all();
    return true;
}

void MuxSocket::join() {
    if (thread_.joinable()) thread_.join();
}

static void put_be64(uint8_t *p, uint64_t v) {
    for (int i = 7; i >= 0; --i) {
        p[i] = static_cast<uint8_t>(v);
        v >>= 8;
    }
}
static uint64_t get_be64(const uint8_t *p) {
    uint64_t v = 0;
    for (int i = 0; i < 8; ++i) v = (v << 8) | p[i];
    return v;
}

std::shared_ptr<SendTicket> MuxSocket::send_bytes(uint64_t tag, uint64_t seq, const void *data, size_t n,
                                                  size_t frame_bytes) {
    auto ticket = std::make_shared<SendTicket>();
    if (!is_open() || mode_ != Mode::Tx) {
        ticket->finish(false);
        return ticket;
    }
    TxItem item{tag, seq, static_cast<const uint8_t *>(data), n, frame_bytes ? frame_bytes : n, {}, ticket};
    if (!enqueue(std::move(item))) ticket->finish(false);
    return ticket;
}

// Blocks (politely) while the ring is full: the TX thread frees a cell with every request it finishes.
bool MuxSocket::enqueue(TxItem &&item) {
    for (int spins = 0; !tx_queue_.try_push(std::move(item)); ++spins) {
        if (!is_open() || stop_) return false;
        if (spins < 64)
            std::this_thread::yield();
        else
            std::this_thread::sleep_for(std::chrono::microseconds(50));
    }
    tx_signal_.bump();
    // The TX thread may have left for good between the is_open() check and the push; then the pusher takes over
    // the consumer role (serialised by the mutex) so that no ticket is left waiting.
    if (tx_exited_.load(std::memory_order_seq_cst)) {
        std::lock_guard<std::mutex> lock(tx_drain_mu_);
        fail_all_tx();
    }
    return true;
}

bool MuxSocket::send_owned(uint64_t tag, uint64_t seq, std::vector<uint8_t> bytes) {
    if (!is_open() || mode_ != Mode::Tx) return false;
    TxItem item{tag, seq, nullptr, bytes.size(), bytes.size(), std::move(bytes), nullptr};
    return enqueue(std::move(item));
}

// the ring's single consumer: the TX thread, or -- after it has exited -- whoever holds tx_drain_mu_
void MuxSocket::fail_all_tx() {
    TxItem it;
    while (tx_queue_.try_pop(it))
        if (it.ticket) it.ticket->finish(false);
}

void MuxSocket::tx_loop() {
    for (;;) {
        TxItem item;
        bool have = false;
        while (!stop_) {
            const uint32_t seen = tx_signal_.load();
            if ((have = tx_queue_.try_pop(item))) break;
            tx_signal_.wait_while(seen, 200); // a push after the failed pop changes the word: no lost wake-up
        }
        if (!have) break;
        const uint8_t *src = item.data ? item.data : item.owned.data();
        size_t left = item.n;
        bool ok = true;
        do { // at least one frame, so zero-length payloads travel too
            const size_t part = left < item.frame_bytes ? left : item.frame_bytes;
            uint8_t head[24];
            put_be64(head, 16 + part);
            put_be64(head + 8, item.tag);
            put_be64(head + 16, item.seq);
            iovec iov[2] = {{head, sizeof head}, {const_cast<uint8_t *>(src), part}};
            msghdr msg{};
            msg.msg_iov = iov;
            msg.msg_iovlen = part ? 2 : 1;
            size_t total = sizeof head + part, sent = 0;
            while (sent < total) {
                const ssize_t k = ::sendmsg(fd_, &msg, MSG_NOSIGNAL);
                if (k < 0) {
                    if (errno == EINTR) continue;
                    ok = false;
                    break;
                }
                sent += static_cast<size_t>(k);
                size_t adv = static_cast<size_t>(k);
                while (adv > 0 && msg.msg_iovlen > 0) {
                    if (adv >= msg.msg_iov[0].iov_len) {
                        adv -= msg.msg_iov[0].iov_len;
                        msg.msg_iov++;
                        msg.msg_iovlen--;
                    } else {
                        msg.msg_iov[0].iov_base = static_cast<uint8_t *>(msg.msg_iov[0].iov_base) + adv;
                        msg.msg_iov[0].iov_len -= adv;
                        adv = 0;
                    }
                }
            }
            src += part;
            left -= part;
        } while (ok && left > 0);
        if (item.ticket) item.ticket->finish(ok);
        if (!ok) {
            open_.store(false, std::memory_order_release);
            break;
        }
    }
    open_.store(false, std::memory_order_release);
    tx_exited_.store(true, std::memory_order_seq_cst); // before the last drain: see enqueue()
    std::lock_guard<std::mutex> lock(tx_drain_mu_);
    fail_all_tx();
}

void MuxSocket::rx_loop() {
    for (;;) {
        uint8_t head[24];
        if (!recv_all(fd_, head, sizeof head)) break;
        const uint64_t len = get_be64(head);
        if (len < 16 || len - 16 > wire::kMaxP2PFrame) {
            PLOG(WARN) << "illegal p2p frame length " << len << " from " << to_string(remote_);
            break;
        }
        RxFrame frame;
        const uint64_t tag = get_be64(head + 8);
        frame.seq = get_be64(head + 16);
        frame.data.n = len - 16;
        frame.data.bytes.reset(new uint8_t[frame.data.n ? frame.data.n : 1]);
        if (frame.data.n && !recv_all(fd_, frame.data.bytes.get(), frame.data.n)) break;
        {
            std::lock_guard<std::mutex> lock(rx_mu_);
            rx_streams_[tag].push_back(std::move(frame));
        }
        rx_cv_.notify_all();
    }
    open_.store(false, std::memory_order_release);
    rx_cv_.notify_all();
}

std::optional<MuxSocket::Frame> MuxSocket::recv_frame(uint64_t tag, uint64_t seq, int timeout_ms) {
    std::unique_lock<std::mutex> lock(rx_mu_);
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(timeout_ms);
    for (;;) {
        auto it = rx_streams_.find(tag);
        if (it != rx_streams_.end()) {
            auto &q = it->second;
            while (!q.empty()) {
                RxFrame f = std::move(q.front());
                q.pop_front();
                if (f.seq != seq) continue; // remains of an aborted operation
                return std::move(f.data);
            }
        }
        if (!is_open() || timeout_ms == 0) return std::nullopt;
        if (rx_cv_.wait_until(lock, deadline) == std::cv_status::timeout) return std::nullopt;
    }
}

void MuxSocket::purge_tag(uint64_t tag) {
    std::lock_guard<std::mutex> lock(rx_mu_);
    rx_streams_.erase(tag);
}

} // namespace pccl::net
