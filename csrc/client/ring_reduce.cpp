#include "ring_reduce.hpp"

#include <algorithm>
#include <cstring>
#include <future>
#include <thread>

#include "../kernels/cpu/reduce.hpp"
#include "../util/env.hpp"
#include "../util/log.hpp"
#include "../util/threadpool.hpp"

namespace pccl::client {

using wire::DType;
using wire::QuantAlgo;
using wire::RedOp;

namespace {

constexpr int kPollMs = 2;
constexpr size_t kOffloadBytes = 256u << 10; // smaller code buffers are quantised inline: a hand-over costs more than it hides

// Helpers that run the sending half of a quantised ring step beside the receiving half.  Shared by all operations of the
// process, never destroyed (its threads sleep on a condition variable when the process exits).  A job only computes and
// enqueues sends; it never waits for another job, so a busy pool delays an operation but cannot block it for good.
ThreadPool &host_kernel_pool() {
    static ThreadPool *pool = [] {
        const unsigned hw = std::max(1u, std::thread::hardware_concurrency());
        const size_t n = static_cast<size_t>(env_u64("PCCL_HOST_KERNEL_THREADS", std::clamp(hw / 4, 1u, 4u)));
        return new ThreadPool(std::max<size_t>(1, n), 1024);
    }();
    return *pool;
}

struct Stage {
    const RingContext &ctx;
    RingResult &res;
    bool failed = false;
    size_t lane; // connection used for metadata and small transfers

    Stage(const RingContext &c, RingResult &r) : ctx(c), res(r), lane(c.tx.empty() ? 0 : c.seq % c.tx.size()) {}

    bool links_open() const {
        for (const auto &s : ctx.tx)
            if (!s->is_open()) return false;
        for (const auto &s : ctx.rx)
            if (!s->is_open()) return false;
        return true;
    }

    bool check_abort() {
        if (ctx.abort_requested && ctx.abort_requested()) {
            res.aborted_by_master = true;
            failed = true;
        }
        if (!links_open()) failed = true;
        return failed;
    }

    // How a transfer of `bytes` is spread over the pool: part i covers [begin_i, end_i).
    size_t parts(size_t bytes) const { return bytes >= ctx.stripe_threshold && ctx.tx.size() > 1 ? ctx.tx.size() : 1; }
    static void part_bounds(size_t bytes, size_t n_parts, size_t i, size_t &begin, size_t &end) {
        // 64-byte aligned cuts keep whole elements of every type together
        const size_t per = ((bytes + n_parts - 1) / n_parts + 63) / 64 * 64;
        begin = std::min(bytes, i * per);
        end = std::min(bytes, begin + per);
    }

    using Tickets = std::vector<std::shared_ptr<net::SendTicket>>;

    Tickets send(const uint8_t *data, size_t bytes) {
        Tickets tickets;
        const size_t n = parts(bytes);
        for (size_t i = 0; i < n; ++i) {
            size_t b, e;
            part_bounds(bytes, n, i, b, e);
            const size_t sock = n == 1 ? lane : i;
            tickets.push_back(ctx.tx[sock]->send_bytes(ctx.tag, ctx.seq, data + b, e - b, ctx.frame_bytes));
        }
        res.tx_bytes += bytes;
        return tickets;
    }

    // Like send(), but the bytes do not exist yet: `produce(b, e)` fills codes[b, e) (quantises that range) right before it
    // is handed to the socket, in pieces, so the first piece is on the wire while the second is being produced.  Same byte
    // stream and the same cuts across the pool as send(): the receiver cannot tell the difference.
    template <typename Produce> Tickets send_pieces(const uint8_t *codes, size_t bytes, Produce produce) {
        constexpr size_t kPiece = 512u << 10;
        Tickets tickets;
        const size_t n = parts(bytes);
        for (size_t i = 0; i < n; ++i) {
            size_t b, e;
            part_bounds(bytes, n, i, b, e);
            net::MuxSocket &sock = *ctx.tx[n == 1 ? lane : i];
            if (b == e) {
                tickets.push_back(sock.send_bytes(ctx.tag, ctx.seq, codes + b, 0, ctx.frame_bytes));
                continue;
            }
            for (size_t p = b; p < e; p += kPiece) {
                const size_t q = std::min(e, p + kPiece);
                produce(p, q);
                tickets.push_back(sock.send_bytes(ctx.tag, ctx.seq, codes + p, q - p, ctx.frame_bytes));
            }
        }
        res.tx_bytes += bytes;
        return tickets;
    }

    // Blocks until every ticket is done, polling for aborts.  After an abort the tickets are still
    // awaited (the successor's RX threads keep draining, or the socket dies) because the TX
    // threads read from caller memory.
    bool finish_send(const Tickets &tickets) {
        bool ok = true;
        for (const auto &ticket : tickets) {
            for (;;) {
                const int st = ticket->wait_for(kPollMs);
                if (st == 1) break;
                if (st == -1) {
                    failed = true;
                    ok = false;
                    break;
                }
                if (check_abort() && res.aborted_by_master) {
                    PLOG(DEBUG) << "abort while a send is in flight, waiting for the TX thread to let go";
                    bool done = false;
                    for (int i = 0; i < 5000 && !done; ++i) done = ticket->wait_for(kPollMs) != 0;
                    if (!done) { // a wedged link is cut so the TX thread lets go of the buffer
                        for (const auto &s : ctx.tx) s->interrupt();
                        ticket->wait_for(1000);
                    }
                    ok = false;
                    break;
                }
            }
        }
        return ok;
    }

    // Receives exactly `bytes` payload bytes, handing every frame to `sink(offset, data, n)`.
    template <typename Sink> bool receive(size_t bytes, Sink sink) {
        const size_t n = parts(bytes);
        for (size_t i = 0; i < n; ++i) {
            size_t b, e;
            part_bounds(bytes, n, i, b, e);
            net::MuxSocket &sock = *ctx.rx[n == 1 ? lane : i];
            size_t got = 0;
            const size_t want = e - b;
            bool first = true;
            while (got < want || (first && want == 0)) {
                auto frame = sock.recv_frame(ctx.tag, ctx.seq, kPollMs);
                if (!frame) {
                    if (check_abort()) return false;
                    continue;
                }
                first = false;
                if (frame->size() > want - got) {
                    PLOG(ERR) << "peer sent more bytes than the chunk holds";
                    failed = true;
                    return false;
                }
                if (frame->size()) sink(b + got, frame->data(), frame->size()); // empty chunks travel as empty frames
                got += frame->size();
                res.rx_bytes += frame->size();
            }
        }
        return true;
    }

    bool exchange_meta(const cpu::QuantParams &mine, cpu::QuantParams &theirs) {
        const auto pkt = cpu::to_wire(mine, ctx.tag);
        if (!ctx.tx[lane]->send_packet(ctx.tag, ctx.seq, pkt)) {
            failed = true;
            return false;
        }
        res.tx_bytes += 8 + 2 + pkt.wire_size();
        for (;;) {
            auto got = ctx.rx[lane]->recv_packet<wire::P2PDequantMeta>(ctx.tag, ctx.seq, kPollMs);
            if (got) {
                res.rx_bytes += 8 + 2 + got->wire_size();
                if (!cpu::from_wire(*got, theirs)) {
                    failed = true;
                    return false;
                }
                return true;
            }
            if (check_abort()) return false;
        }
    }
};

} // namespace

RingResult ring_allreduce(const RingContext &ctx, const ReduceArgs &a, std::vector<uint8_t> *backup) {
    RingResult res;
    const size_t esz = wire::dtype_size(a.dtype);
    auto *dst = static_cast<uint8_t *>(a.dst);
    const bool in_place = a.src == a.dst;
    if (backup) backup->clear();

    if (ctx.world < 2) {
        if (!in_place) std::memcpy(dst, a.src, a.count * esz);
        res.ok = true;
        return res;
    }
    if (ctx.tx.empty() || ctx.rx.empty() || ctx.tx.size() != ctx.rx.size()) return res;

    // In-place operation destroys the input while it runs; keep a copy to roll back to.  Out of place nothing is copied up
    // front: in the reduce-scatter every chunk but the own one is received exactly once, so the input is brought over piece
    // by piece right before the incoming frame is folded into it (cache-warm, and while the network is busy anyway); the
    // first chunk sent is read from the input, and the all-gather overwrites whatever was not reduced here.
    const auto *src = static_cast<const uint8_t *>(a.src);
    if (in_place && backup) backup->assign(dst, dst + a.count * esz);
    auto fail = [&]() {
        PLOG(DEBUG) << "ring all-reduce tag " << ctx.tag << " seq " << ctx.seq << " failed locally (aborted_by_master=" << res.aborted_by_master << ")";
        if (in_place && backup && !backup->empty()) std::memcpy(dst, backup->data(), backup->size());
        res.ok = false;
        return res;
    };

    const bool quantised = a.quant != QuantAlgo::None;
    const bool block_format = a.quant == QuantAlgo::BlockMinMax || a.quant == QuantAlgo::BlockFp8;
    cpu::QuantParams qp;
    qp.algo = a.quant;
    qp.value_dtype = a.dtype;
    qp.code_dtype = a.quant_dtype;
    const size_t qsz = quantised && !block_format ? wire::dtype_size(a.quant_dtype) : 1;

    Stage st(ctx, res);
    const uint32_t W = ctx.world, r = ctx.rank;
    std::vector<uint8_t> tx_codes, rx_codes;

    // ---------------- reduce-scatter
    for (uint32_t step = 0; step + 1 < W; ++step) {
        const uint32_t tx_chunk = (r + W - step) % W, rx_chunk = (r + W - step - 1) % W;
        uint64_t tb, te, rb, re;
        chunk_bounds(a.count, W, tx_chunk, tb, te);
        chunk_bounds(a.count, W, rx_chunk, rb, re);
        const uint64_t tx_n = te - tb, rx_n = re - rb;

        // what this peer sends first is its own, still untouched chunk; later it forwards what it reduced into dst
        const uint8_t *tx_from = (in_place || step > 0 ? dst : src) + tb * esz;
        if (!quantised) {
            const auto tickets = st.send(tx_from, tx_n * esz);
            const bool rx_ok = st.receive(rx_n * esz, [&](size_t off, const uint8_t *data, size_t n) {
                if (!in_place) std::memcpy(dst + rb * esz + off, src + rb * esz + off, n);
                cpu::reduce_inplace(dst + rb * esz + off, data, n / esz, a.dtype, a.op);
            });
            const bool tx_ok = st.finish_send(tickets);
            if (!rx_ok || !tx_ok) return fail();
            continue;
        }

        // The sending half (quantise piece by piece, hand each piece to the sockets) runs beside the receiving half
        // (decode and reduce every frame as it arrives): the two touch different chunks of dst.
        cpu::QuantParams mine = qp, theirs = qp;
        cpu::quant_analyze(tx_from, tx_n, mine);
        if (!st.exchange_meta(mine, theirs)) return fail();
        tx_codes.resize(cpu::quantized_size(mine, tx_n));
        Stage::Tickets tickets;
        auto tx_half = [&] {
            if (block_format) { // records first, codes after them: produced as a whole
                cpu::quantize(tx_from, tx_codes.data(), tx_n, mine);
                tickets = st.send(tx_codes.data(), tx_codes.size());
            } else {
                tickets = st.send_pieces(tx_codes.data(), tx_codes.size(), [&](size_t b, size_t e) {
                    cpu::quantize(tx_from + (b / qsz) * esz, tx_codes.data() + b, (e - b) / qsz, mine);
                });
            }
        };
        std::future<void> tx_job;
        if (tx_codes.size() >= kOffloadBytes) tx_job = host_kernel_pool().submit(tx_half);
        else tx_half();
        const size_t rx_bytes = cpu::quantized_size(theirs, rx_n);
        bool rx_ok;
        if (block_format) {
            rx_codes.resize(rx_bytes);
            rx_ok = st.receive(rx_bytes, [&](size_t off, const uint8_t *data, size_t n) { std::memcpy(rx_codes.data() + off, data, n); });
            if (rx_ok && !in_place) std::memcpy(dst + rb * esz, src + rb * esz, rx_n * esz);
            if (rx_ok) cpu::dequantize_reduce(dst + rb * esz, rx_codes.data(), rx_n, theirs, a.op);
        } else {
            rx_ok = st.receive(rx_bytes, [&](size_t off, const uint8_t *data, size_t n) {
                if (!in_place) std::memcpy(dst + (rb + off / qsz) * esz, src + (rb + off / qsz) * esz, n / qsz * esz);
                cpu::dequantize_reduce(dst + (rb + off / qsz) * esz, data, n / qsz, theirs, a.op);
            });
        }
        if (tx_job.valid()) tx_job.get(); // before anything else: it works on this frame's variables
        const bool tx_ok = st.finish_send(tickets);
        if (!rx_ok || !tx_ok) return fail();
    }

    // ---------------- all-gather
    cpu::QuantParams forwarded = qp; // metadata that travels with the codes we forward
    for (uint32_t step = 0; step + 1 < W; ++step) {
        const uint32_t tx_chunk = (r + 1 + W - step) % W, rx_chunk = (r + W - step) % W;
        uint64_t tb, te, rb, re;
        chunk_bounds(a.count, W, tx_chunk, tb, te);
        chunk_bounds(a.count, W, rx_chunk, rb, re);
        const uint64_t tx_n = te - tb, rx_n = re - rb;

        if (!quantised) {
            const auto tickets = st.send(dst + tb * esz, tx_n * esz);
            const bool rx_ok = st.receive(rx_n * esz, [&](size_t off, const uint8_t *data, size_t n) {
                std::memcpy(dst + rb * esz + off, data, n);
            });
            const bool tx_ok = st.finish_send(tickets);
            if (!rx_ok || !tx_ok) return fail();
            continue;
        }

        const bool own = step == 0;
        if (own) {
            // The chunk this peer owns is quantised exactly once, and the owner adopts the decoded values, so it holds what
            // everybody else will decode.  Later steps forward received codes and metadata unchanged.
            forwarded = qp;
            cpu::quant_analyze(dst + tb * esz, tx_n, forwarded);
            tx_codes.resize(cpu::quantized_size(forwarded, tx_n));
        }
        cpu::QuantParams theirs = qp;
        if (!st.exchange_meta(forwarded, theirs)) return fail();
        Stage::Tickets tickets;
        auto tx_half = [&] {
            if (!own) {
                tickets = st.send(tx_codes.data(), tx_codes.size());
            } else if (block_format) {
                cpu::quantize(dst + tb * esz, tx_codes.data(), tx_n, forwarded);
                tickets = st.send(tx_codes.data(), tx_codes.size());
                cpu::dequantize_reduce(dst + tb * esz, tx_codes.data(), tx_n, forwarded, RedOp::Set);
            } else {
                tickets = st.send_pieces(tx_codes.data(), tx_codes.size(), [&](size_t b, size_t e) {
                    cpu::quantize(dst + (tb + b / qsz) * esz, tx_codes.data() + b, (e - b) / qsz, forwarded);
                    cpu::dequantize_reduce(dst + (tb + b / qsz) * esz, tx_codes.data() + b, (e - b) / qsz, forwarded, RedOp::Set);
                });
            }
        };
        std::future<void> tx_job;
        if (own && tx_codes.size() >= kOffloadBytes) tx_job = host_kernel_pool().submit(tx_half);
        else tx_half();
        const size_t rx_bytes = cpu::quantized_size(theirs, rx_n);
        rx_codes.resize(rx_bytes);
        const bool rx_ok = st.receive(rx_bytes, [&](size_t off, const uint8_t *data, size_t n) {
            std::memcpy(rx_codes.data() + off, data, n);
            if (!block_format) cpu::dequantize_reduce(dst + (rb + off / qsz) * esz, data, n / qsz, theirs, RedOp::Set); // while it is in cache
        });
        if (tx_job.valid()) tx_job.get();
        const bool tx_ok = st.finish_send(tickets);
        if (!rx_ok || !tx_ok) return fail();
        if (block_format) cpu::dequantize_reduce(dst + rb * esz, rx_codes.data(), rx_n, theirs, RedOp::Set);
        // next step forwards the received codes and metadata unchanged (no second quantisation)
        tx_codes.swap(rx_codes);
        forwarded = theirs;
    }

    if (a.op == RedOp::Avg) cpu::finalize_avg(dst, a.count, a.dtype, W);
    res.ok = true;
    return res;
}

// a "local ring" of one peer reduces nothing: the input then has to be in place before the ring between hosts starts
static bool L_is_one(const RingContext &local) { return local.world <= 1; }

RingResult hierarchical_allreduce(const RingContext &local, const RingContext &cross, const ReduceArgs &a, std::vector<uint8_t> *backup) {
    RingResult res;
    const size_t esz = wire::dtype_size(a.dtype);
    auto *dst = static_cast<uint8_t *>(a.dst);
    const bool in_place = a.src == a.dst;
    if (backup) backup->clear();
    if (local.world < 1 || cross.world < 1) return res;
    const uint32_t L = local.world, B = cross.world, l = local.rank;
    if ((L > 1 && (local.tx.empty() || local.tx.size() != local.rx.size())) || (B > 1 && (cross.tx.empty() || cross.tx.size() != cross.rx.size())))
        return res;
    // (out of place the input is brought over piece by piece as the frames of the local reduce-scatter arrive, as in the
    // flat ring: every chunk but the own one is received exactly once there, and the all-gather overwrites the rest)
    const auto *src = static_cast<const uint8_t *>(a.src);
    if (in_place && backup) backup->assign(dst, dst + a.count * esz);
    if (!in_place && L_is_one(local)) std::memcpy(dst, a.src, a.count * esz);
    auto fail = [&]() {
        if (in_place && backup && !backup->empty()) std::memcpy(dst, backup->data(), backup->size());
        res.ok = false;
        return res;
    };
    // AVG is a sum that is divided by the number of ALL peers at the very end
    const RedOp op = a.op == RedOp::Avg ? RedOp::Sum : a.op;

    // ---- 1. reduce-scatter among the peers of this host (same schedule as the flat ring)
    Stage st(local, res);
    for (uint32_t step = 0; step + 1 < L; ++step) {
        const uint32_t tx_chunk = (l + L - step) % L, rx_chunk = (l + L - step - 1) % L;
        uint64_t tb, te, rb, re;
        chunk_bounds(a.count, L, tx_chunk, tb, te);
        chunk_bounds(a.count, L, rx_chunk, rb, re);
        const auto tickets = st.send((in_place || step > 0 ? dst : src) + tb * esz, (te - tb) * esz);
        const bool rx_ok = st.receive((re - rb) * esz, [&](size_t off, const uint8_t *data, size_t n) {
            if (!in_place) std::memcpy(dst + rb * esz + off, src + rb * esz + off, n);
            cpu::reduce_inplace(dst + rb * esz + off, data, n / esz, a.dtype, op);
        });
        const bool tx_ok = st.finish_send(tickets);
        if (!rx_ok || !tx_ok) return fail();
    }

    // ---- 2. the chunk this peer now owns, (l+1) mod L, is summed over the hosts by the peers of equal local rank.  A wire
    //         quantisation applies here and only here: between hosts is where bytes are expensive.  The owner of every piece
    //         adopts what the others decode (ring_allreduce), the local all-gather then copies exact bytes, so all L*B peers
    //         end with identical bits.
    {
        uint64_t ob, oe;
        chunk_bounds(a.count, L, (l + 1) % L, ob, oe);
        ReduceArgs sub = a;
        sub.src = sub.dst = dst + ob * esz;
        sub.count = oe - ob;
        sub.op = op;
        RingContext cx = cross;
        cx.tag = local.tag;
        cx.seq = local.seq;
        cx.abort_requested = local.abort_requested;
        const RingResult r = ring_allreduce(cx, sub, nullptr);
        res.tx_bytes += r.tx_bytes;
        res.rx_bytes += r.rx_bytes;
        res.aborted_by_master = res.aborted_by_master || r.aborted_by_master;
        if (!r.ok) return fail();
    }

    // ---- 3. all-gather among the peers of this host
    for (uint32_t step = 0; step + 1 < L; ++step) {
        const uint32_t tx_chunk = (l + 1 + L - step) % L, rx_chunk = (l + L - step) % L;
        uint64_t tb, te, rb, re;
        chunk_bounds(a.count, L, tx_chunk, tb, te);
        chunk_bounds(a.count, L, rx_chunk, rb, re);
        const auto tickets = st.send(dst + tb * esz, (te - tb) * esz);
        const bool rx_ok = st.receive((re - rb) * esz, [&](size_t off, const uint8_t *data, size_t n) { std::memcpy(dst + rb * esz + off, data, n); });
        const bool tx_ok = st.finish_send(tickets);
        if (!rx_ok || !tx_ok) return fail();
    }
    if (a.op == RedOp::Avg) cpu::finalize_avg(dst, a.count, a.dtype, L * B);
    res.ok = true;
    return res;
}

} // namespace pccl::client
