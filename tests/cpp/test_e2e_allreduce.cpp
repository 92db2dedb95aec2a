// End-to-end all-reduce contracts (reference ccoip/tests/end_to_end/test_all_reduce.cpp).
#include <cmath>
#include <cstring>
#include <numeric>

#include "e2e_util.hpp"
#include "net/blocking_socket.hpp"
#include "util/guarded_alloc.hpp"

using namespace pccl;
using client::ReduceOutcome;
using client::ReduceRequest;

template <typename T> constexpr wire::DType dtype_of();
template <> constexpr wire::DType dtype_of<uint8_t>() { return wire::DType::U8; }
template <> constexpr wire::DType dtype_of<int8_t>() { return wire::DType::I8; }
template <> constexpr wire::DType dtype_of<uint16_t>() { return wire::DType::U16; }
template <> constexpr wire::DType dtype_of<int16_t>() { return wire::DType::I16; }
template <> constexpr wire::DType dtype_of<uint32_t>() { return wire::DType::U32; }
template <> constexpr wire::DType dtype_of<int32_t>() { return wire::DType::I32; }
template <> constexpr wire::DType dtype_of<uint64_t>() { return wire::DType::U64; }
template <> constexpr wire::DType dtype_of<int64_t>() { return wire::DType::I64; }
template <> constexpr wire::DType dtype_of<float>() { return wire::DType::F32; }
template <> constexpr wire::DType dtype_of<double>() { return wire::DType::F64; }

template <typename T> static T fold(T a, T b, wire::RedOp op) {
    switch (op) {
    case wire::RedOp::Sum:
    case wire::RedOp::Avg: return static_cast<T>(a + b);
    case wire::RedOp::Prod: return static_cast<T>(a * b);
    case wire::RedOp::Max: return a > b ? a : b;
    case wire::RedOp::Min: return a < b ? a : b;
    default: return b;
    }
}

template <typename T> static void run_case(int world, size_t count, wire::RedOp op, bool in_place) {
    TestMaster tm;
    auto peers = connect_peers(tm, world);
    std::vector<std::vector<T>> inputs(world, std::vector<T>(count)), outputs(world, std::vector<T>(count));
    for (int r = 0; r < world; ++r)
        for (size_t i = 0; i < count; ++i) inputs[r][i] = static_cast<T>((i * 7 + r * 3 + 1) % 11 + (op == wire::RedOp::Prod ? 0 : r));
    std::vector<T> expect(inputs[0]);
    for (int r = 1; r < world; ++r)
        for (size_t i = 0; i < count; ++i) expect[i] = fold(expect[i], inputs[r][i], op);
    if (op == wire::RedOp::Avg)
        for (auto &v : expect) v = static_cast<T>(v / static_cast<T>(world));

    std::vector<std::thread> ts;
    for (int r = 0; r < world; ++r) {
        ts.emplace_back([&, r] {
            ReduceRequest req;
            std::vector<T> work = inputs[r];
            req.send = in_place ? work.data() : inputs[r].data();
            req.recv = in_place ? work.data() : outputs[r].data();
            req.count = count;
            req.dtype = dtype_of<T>();
            req.op = op;
            req.tag = 1;
            CHECK(peers[r]->all_reduce_async(req) == client::Client::Launch::Ok);
            ReduceOutcome out;
            CHECK(peers[r]->await_reduce(1, out));
            CHECK_EQ(out.world, static_cast<uint32_t>(world));
            if (in_place) outputs[r] = work;
        });
    }
    for (auto &t : ts) t.join();
    for (int r = 0; r < world; ++r)
        for (size_t i = 0; i < count; ++i) {
            if constexpr (std::is_floating_point_v<T>)
                CHECK(std::fabs(static_cast<double>(outputs[r][i]) - static_cast<double>(expect[i])) < 1e-5);
            else
                CHECK(outputs[r][i] == expect[i]);
        }
}

template <typename T> static void typed_suite() {
    run_case<T>(2, 203530, wire::RedOp::Sum, false);
    run_case<T>(2, 203530, wire::RedOp::Sum, true);
    for (size_t n : {size_t(2), size_t(3), size_t(1025)}) run_case<T>(3, n, wire::RedOp::Sum, false);
    run_case<T>(4, 2, wire::RedOp::Sum, true); // some ranks own empty chunks
    run_case<T>(2, 1000, wire::RedOp::Avg, false);
    run_case<T>(3, 1000, wire::RedOp::Min, false);
    run_case<T>(3, 1000, wire::RedOp::Max, true);
}

TEST(allreduce_u8) { typed_suite<uint8_t>(); }
TEST(allreduce_i16) { typed_suite<int16_t>(); }
TEST(allreduce_u32) { typed_suite<uint32_t>(); }
TEST(allreduce_i64) { typed_suite<int64_t>(); }
TEST(allreduce_f32) { typed_suite<float>(); }
TEST(allreduce_f64) { typed_suite<double>(); }
TEST(allreduce_other_ints) {
    run_case<int8_t>(3, 1025, wire::RedOp::Sum, false);
    run_case<uint16_t>(3, 1025, wire::RedOp::Max, false);
    run_case<int32_t>(4, 4099, wire::RedOp::Sum, true);
    run_case<uint64_t>(2, 77, wire::RedOp::Prod, false);
}

// 42 + 43 must come out as exactly 85 through uint8 zero-point/scale quantisation
template <typename T> static void quantised_exact() {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::vector<std::thread> ts;
    for (int r = 0; r < 2; ++r) {
        ts.emplace_back([&, r] {
            std::vector<T> in(1024, static_cast<T>(42 + r)), out(1024, 0);
            ReduceRequest req;
            req.send = in.data();
            req.recv = out.data();
            req.count = in.size();
            req.dtype = dtype_of<T>();
            req.op = wire::RedOp::Sum;
            req.quant = wire::QuantAlgo::ZeroPointScale;
            req.quant_dtype = wire::DType::U8;
            req.tag = 1;
            CHECK(peers[r]->all_reduce_async(req) == client::Client::Launch::Ok);
            ReduceOutcome o;
            CHECK(peers[r]->await_reduce(1, o));
            for (auto v : out) CHECK(v == static_cast<T>(85));
        });
    }
    for (auto &t : ts) t.join();
}
TEST(quantised_zps_exact_f32) { quantised_exact<float>(); }
TEST(quantised_zps_exact_f64) { quantised_exact<double>(); }

// min-max uint8 quantisation: every peer ends up with bit-identical results close to the truth
TEST(quantised_minmax_bit_identical) {
    for (int world : {2, 3, 4}) {
        TestMaster tm;
        auto peers = connect_peers(tm, world);
        const size_t n = 10007;
        std::vector<std::vector<float>> in(world, std::vector<float>(n)), out(world, std::vector<float>(n));
        for (int r = 0; r < world; ++r)
            for (size_t i = 0; i < n; ++i) in[r][i] = std::sin(0.001f * i + r) * 3.0f;
        std::vector<std::thread> ts;
        for (int r = 0; r < world; ++r)
            ts.emplace_back([&, r] {
                ReduceRequest req;
                req.send = in[r].data();
                req.recv = out[r].data();
                req.count = n;
                req.op = wire::RedOp::Sum;
                req.quant = wire::QuantAlgo::MinMax;
                req.quant_dtype = wire::DType::U8;
                req.tag = 7;
                CHECK(peers[r]->all_reduce_async(req) == client::Client::Launch::Ok);
                ReduceOutcome o;
                CHECK(peers[r]->await_reduce(7, o));
                CHECK(o.tx_bytes > 0 && o.tx_bytes < n * sizeof(float) * 2); // less than unquantised traffic
            });
        for (auto &t : ts) t.join();
        for (int r = 1; r < world; ++r) CHECK(std::memcmp(out[0].data(), out[r].data(), n * sizeof(float)) == 0);
        for (size_t i = 0; i < n; ++i) {
            float truth = 0;
            for (int r = 0; r < world; ++r) truth += in[r][i];
            CHECK(std::fabs(out[0][i] - truth) < 0.05f * world);
        }
    }
}

// Large quantised operands: the sending half of every ring step (quantise in pieces, send each piece) runs on a helper beside
// the receiving half, over a striped connection pool.  Whatever the interleaving, the result is a function of the inputs
// alone: all peers agree bit for bit, and a second run gives the same bits as the first.
TEST(large_quantised_operands_are_deterministic) {
    TestMaster tm;
    const int world = 3;
    std::vector<std::unique_ptr<client::Client>> peers;
    {
        auto built = connect_peers(tm, world);
        peers = std::move(built);
    }
    const size_t n = 3'000'017;
    std::vector<std::vector<float>> in(world, std::vector<float>(n));
    for (int r = 0; r < world; ++r)
        for (size_t i = 0; i < n; ++i) in[r][i] = std::sin(0.0007f * static_cast<float>(i % 100'000) + r) * (1.0f + static_cast<float>(i % 7));
    struct Case {
        wire::QuantAlgo algo;
        wire::DType code;
        float tol;
    };
    for (const Case c : {Case{wire::QuantAlgo::MinMax, wire::DType::U8, 0.25f}, Case{wire::QuantAlgo::ZeroPointScale, wire::DType::U8, 0.25f},
                         Case{wire::QuantAlgo::MinMax, wire::DType::U16, 0.002f}, Case{wire::QuantAlgo::BlockMinMax, wire::DType::U8, 0.25f},
                         Case{wire::QuantAlgo::BlockFp8, wire::DType::F8E4M3, 1.2f}}) {
        std::vector<std::vector<float>> first;
        for (int run = 0; run < 2; ++run) {
            std::vector<std::vector<float>> out(world);
            std::vector<std::thread> ts;
            for (int r = 0; r < world; ++r)
                ts.emplace_back([&, r] {
                    out[r] = in[r]; // in place
                    ReduceRequest req;
                    req.send = req.recv = out[r].data();
                    req.count = n;
                    req.op = wire::RedOp::Avg;
                    req.quant = c.algo;
                    req.quant_dtype = c.code;
                    req.tag = 11;
                    ReduceOutcome o;
                    CHECK(peers[r]->all_reduce_sync(req, o) == client::Client::Launch::Ok);
                    CHECK(o.ok);
                });
            for (auto &t : ts) t.join();
            for (int r = 1; r < world; ++r) CHECK(std::memcmp(out[0].data(), out[r].data(), n * sizeof(float)) == 0);
            for (size_t i = 0; i < n; i += 13) {
                const float truth = (in[0][i] + in[1][i] + in[2][i]) / 3.0f;
                CHECK(std::fabs(out[0][i] - truth) < c.tol);
            }
            if (run == 0) first = out;
            else CHECK(std::memcmp(first[0].data(), out[0].data(), n * sizeof(float)) == 0);
        }
    }
}

TEST(quantised_block_formats) {
    for (auto algo : {wire::QuantAlgo::BlockMinMax, wire::QuantAlgo::BlockFp8}) {
        TestMaster tm;
        auto peers = connect_peers(tm, 3);
        const size_t n = 5000;
        std::vector<std::vector<float>> in(3, std::vector<float>(n)), out(3, std::vector<float>(n));
        for (int r = 0; r < 3; ++r)
            for (size_t i = 0; i < n; ++i) in[r][i] = std::cos(0.01f * i * (r + 1)) * (1.0f + 0.001f * i);
        std::vector<std::thread> ts;
        for (int r = 0; r < 3; ++r)
            ts.emplace_back([&, r] {
                ReduceRequest req;
                req.send = in[r].data();
                req.recv = out[r].data();
                req.count = n;
                req.op = wire::RedOp::Avg;
                req.quant = algo;
                req.quant_dtype = algo == wire::QuantAlgo::BlockFp8 ? wire::DType::F8E4M3 : wire::DType::U8;
                req.tag = 3;
                CHECK(peers[r]->all_reduce_async(req) == client::Client::Launch::Ok);
                ReduceOutcome o;
                CHECK(peers[r]->await_reduce(3, o));
            });
        for (auto &t : ts) t.join();
        for (int r = 1; r < 3; ++r) CHECK(std::memcmp(out[0].data(), out[r].data(), n * sizeof(float)) == 0);
        for (size_t i = 0; i < n; ++i) {
            const float truth = (in[0][i] + in[1][i] + in[2][i]) / 3.0f;
            // uint8 blocks resolve 1/255 of the block range; e4m3 keeps 3 mantissa bits (about 6%
            // of the block maximum per hop)
            const float tol = algo == wire::QuantAlgo::BlockFp8 ? 0.9f : 0.08f;
            CHECK(std::fabs(out[0][i] - truth) < tol);
        }
    }
}

// two tags run concurrently with different operators
TEST(concurrent_tags) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    const size_t n = 1 << 16;
    std::vector<std::thread> ts;
    for (int r = 0; r < 2; ++r)
        ts.emplace_back([&, r] {
            std::vector<float> a(n, 1.0f + r), b(n, 10.0f * (r + 1)), ra(n), rb(n);
            ReduceRequest q1, q2;
            q1.send = a.data(), q1.recv = ra.data(), q1.count = n, q1.op = wire::RedOp::Sum, q1.tag = 1;
            q2.send = b.data(), q2.recv = rb.data(), q2.count = n, q2.op = wire::RedOp::Max, q2.tag = 2;
            CHECK(peers[r]->all_reduce_async(q1) == client::Client::Launch::Ok);
            CHECK(peers[r]->all_reduce_async(q2) == client::Client::Launch::Ok);
            // joint calls are refused while reductions are in flight (reference test_all_reduce.cpp:449-524, :527-618)
            CHECK(!peers[r]->update_topology());
            CHECK(!peers[r]->optimize_topology());
            client::SharedStateArgs state;
            client::SharedEntry entry;
            entry.key = "w", entry.data = a.data(), entry.count = n, entry.dtype = wire::DType::F32;
            state.entries.push_back(entry);
            uint64_t tx = 0, rx = 0;
            CHECK(!peers[r]->sync_shared_state(state, tx, rx));
            ReduceOutcome o;
            CHECK(peers[r]->await_reduce(1, o));
            CHECK(peers[r]->await_reduce(2, o));
            for (size_t i = 0; i < n; ++i) {
                CHECK(ra[i] == 3.0f);
                CHECK(rb[i] == 20.0f);
            }
        });
    for (auto &t : ts) t.join();
}

// a tag may only be in flight once; the first operation is unaffected
TEST(duplicate_tag_rejected) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    const size_t n = 1 << 18;
    std::vector<std::thread> ts;
    for (int r = 0; r < 2; ++r)
        ts.emplace_back([&, r] {
            std::vector<float> a(n, 1.0f), ra(n), rb(n);
            ReduceRequest q;
            q.send = a.data(), q.recv = ra.data(), q.count = n, q.tag = 5;
            CHECK(peers[r]->all_reduce_async(q) == client::Client::Launch::Ok);
            ReduceRequest dup = q;
            dup.recv = rb.data();
            CHECK(peers[r]->all_reduce_async(dup) == client::Client::Launch::DuplicateTag);
            ReduceOutcome o;
            CHECK(peers[r]->await_reduce(5, o));
            for (auto v : ra) CHECK(v == 2.0f);
            // tag can be reused afterwards
            CHECK(peers[r]->all_reduce_async(q) == client::Client::Launch::Ok);
            CHECK(peers[r]->await_reduce(5, o));
        });
    for (auto &t : ts) t.join();
}

// basic_reduce_test of BASELINE.json: 2 peers + master on localhost, fp32 sum of 4 elements
TEST(basic_reduce_four_elements) {
    TestMaster tm;
    auto peers = connect_peers(tm, 2);
    std::vector<std::thread> ts;
    for (int r = 0; r < 2; ++r)
        ts.emplace_back([&, r] {
            float in[4] = {1, 2, 3, 4}, out[4] = {0};
            for (int it = 0; it < 50; ++it) {
                ReduceRequest q;
                q.send = in, q.recv = out, q.count = 4, q.tag = 0;
                CHECK(peers[r]->all_reduce_async(q) == client::Client::Launch::Ok);
                ReduceOutcome o;
                CHECK(peers[r]->await_reduce(0, o));
                for (int i = 0; i < 4; ++i) CHECK(out[i] == 2.0f * (i + 1));
            }
        });
    for (auto &t : ts) t.join();
}

// connection pool of 4: large transfers are striped over all connections, small ones use one
TEST(striped_over_connection_pool) {
    TestMaster tm;
    tm.pool_size = 4;
    auto peers = connect_peers(tm, 3);
    for (size_t n : {size_t(5), size_t(3'000'017)}) {
        std::vector<std::vector<float>> in(3, std::vector<float>(n)), out(3, std::vector<float>(n));
        for (int r = 0; r < 3; ++r)
            for (size_t i = 0; i < n; ++i) in[r][i] = static_cast<float>((i * 13 + r) % 97);
        std::vector<std::thread> ts;
        for (int r = 0; r < 3; ++r)
            ts.emplace_back([&, r] {
                for (int rep = 0; rep < 3; ++rep) {
                    ReduceRequest q;
                    q.send = in[r].data(), q.recv = out[r].data(), q.count = n, q.op = wire::RedOp::Sum, q.tag = 9;
                    CHECK(peers[r]->all_reduce_async(q) == client::Client::Launch::Ok);
                    ReduceOutcome o;
                    CHECK(peers[r]->await_reduce(9, o));
                }
            });
        for (auto &t : ts) t.join();
        for (int r = 0; r < 3; ++r)
            for (size_t i = 0; i < n; ++i) CHECK(out[r][i] == in[0][i] + in[1][i] + in[2][i]);
    }
}

// User buffers end exactly at an inaccessible page: any read or write past `count` elements in the reduce,
// quantise or receive paths faults here.  Counts are chosen to leave ragged chunks and ragged quant blocks.
TEST(buffers_in_guarded_blocks) {
    TestMaster tm;
    auto peers = connect_peers(tm, 3);
    uint64_t tag = 10;
    for (auto algo : {wire::QuantAlgo::None, wire::QuantAlgo::MinMax, wire::QuantAlgo::ZeroPointScale, wire::QuantAlgo::BlockMinMax,
                      wire::QuantAlgo::BlockFp8}) {
        for (size_t n : {1u, 2u, 127u, 129u, 1000u, 4099u}) {
            ++tag;
            std::vector<float *> in(3), out(3);
            for (int r = 0; r < 3; ++r) {
                in[r] = static_cast<float *>(guarded_alloc(n * sizeof(float)));
                out[r] = static_cast<float *>(guarded_alloc(n * sizeof(float)));
                for (size_t i = 0; i < n; ++i) in[r][i] = static_cast<float>((i + r) % 5);
            }
            std::vector<std::thread> ts;
            for (int r = 0; r < 3; ++r)
                ts.emplace_back([&, r] {
                    ReduceRequest req;
                    req.send = in[r];
                    req.recv = out[r];
                    req.count = n;
                    req.op = wire::RedOp::Sum;
                    req.quant = algo;
                    req.quant_dtype = algo == wire::QuantAlgo::BlockFp8 ? wire::DType::F8E4M3 : wire::DType::U8;
                    req.tag = tag;
                    CHECK(peers[r]->all_reduce_async(req) == client::Client::Launch::Ok);
                    ReduceOutcome o;
                    CHECK(peers[r]->await_reduce(tag, o));
                });
            for (auto &t : ts) t.join();
            for (size_t i = 0; i < n; ++i) {
                const float truth = static_cast<float>(i % 5 + (i + 1) % 5 + (i + 2) % 5);
                CHECK(std::fabs(out[0][i] - truth) < (algo == wire::QuantAlgo::None ? 1e-6f : 0.8f));
            }
            for (int r = 0; r < 3; ++r) {
                CHECK(guarded_free(in[r]));
                CHECK(guarded_free(out[r]));
            }
        }
    }
}

// DiLoCo outer step on host tensors (two-pass form over the TCP ring): two consecutive steps against a direct
// evaluation of  g = mean(params - inner);  m = first ? g : mu*m + g;  params -= lr*(g + mu*m);  inner = params.
TEST(outer_step_host_tensors) {
    TestMaster tm;
    const int world = 3;
    auto peers = connect_peers(tm, world);
    const size_t n = 1237;
    const float lr = 0.7f, mu = 0.9f;
    std::vector<std::vector<float>> params(world, std::vector<float>(n)), inner(world, std::vector<float>(n)), mom(world, std::vector<float>(n, 123.0f));
    std::vector<float> ref_p(n), ref_m(n, 0.0f);
    for (size_t i = 0; i < n; ++i) ref_p[i] = std::sin(0.01f * i);
    for (int r = 0; r < world; ++r) params[r] = ref_p;
    for (int step = 0; step < 2; ++step) {
        // every peer drifted differently during its inner steps
        for (int r = 0; r < world; ++r)
            for (size_t i = 0; i < n; ++i) inner[r][i] = params[r][i] - 0.001f * static_cast<float>((i + 7 * r + step) % 13);
        for (size_t i = 0; i < n; ++i) {
            float g = 0;
            for (int r = 0; r < world; ++r) g += ref_p[i] - inner[r][i];
            g /= world;
            ref_m[i] = step == 0 ? g : mu * ref_m[i] + g;
            ref_p[i] -= lr * (g + mu * ref_m[i]);
        }
        std::vector<std::thread> ts;
        for (int r = 0; r < world; ++r)
            ts.emplace_back([&, r] {
                device::OuterStepArgs o;
                o.params = params[r].data();
                o.inner = inner[r].data();
                o.momentum = mom[r].data();
                o.lr = lr;
                o.mu = mu;
                o.first = step == 0;
                ReduceRequest req;
                req.count = n;
                req.op = wire::RedOp::Avg;
                req.tag = 40;
                req.outer = &o;
                ReduceOutcome out;
                CHECK(peers[r]->all_reduce_sync(req, out) == client::Client::Launch::Ok);
                CHECK(out.ok);
            });
        for (auto &t : ts) t.join();
        for (int r = 0; r < world; ++r)
            for (size_t i = 0; i < n; ++i) {
                CHECK(std::fabs(params[r][i] - ref_p[i]) < 1e-5f);
                CHECK(std::fabs(mom[r][i] - ref_m[i]) < 1e-5f);
                CHECK(inner[r][i] == params[r][i]);
                CHECK(params[r][i] == params[0][i]); // bit-identical on all peers
            }
    }
}

// ---- two-level ring for groups that span hosts; the hosts are emulated by giving every client a host id -----------------
namespace {
std::vector<std::unique_ptr<client::Client>> connect_on_hosts(const TestMaster &tm, const std::vector<uint64_t> &host_of) {
    return connect_peers(tm, static_cast<int>(host_of.size()), {}, host_of);
}

template <typename T>
void reduce_everywhere(std::vector<std::unique_ptr<client::Client>> &peers, size_t count, wire::RedOp op, bool in_place, pcclExtDataPlane_t want_plane) {
    const int world = static_cast<int>(peers.size());
    std::vector<std::vector<T>> inputs(world, std::vector<T>(count)), outputs(world, std::vector<T>(count));
    for (int r = 0; r < world; ++r)
        for (size_t i = 0; i < count; ++i) inputs[r][i] = static_cast<T>((i * 7 + r * 3 + 1) % 11 + r);
    std::vector<T> expect(inputs[0]);
    for (int r = 1; r < world; ++r)
        for (size_t i = 0; i < count; ++i) expect[i] = fold(expect[i], inputs[r][i], op);
    if (op == wire::RedOp::Avg)
        for (auto &v : expect) v = static_cast<T>(v / static_cast<T>(world));
    std::vector<std::thread> ts;
    for (int r = 0; r < world; ++r)
        ts.emplace_back([&, r] {
            ReduceRequest req;
            std::vector<T> work = inputs[r];
            req.send = in_place ? work.data() : inputs[r].data();
            req.recv = in_place ? work.data() : outputs[r].data();
            req.count = count;
            req.dtype = dtype_of<T>();
            req.op = op;
            req.tag = 5;
            ReduceOutcome out;
            CHECK(peers[r]->all_reduce_sync(req, out) == client::Client::Launch::Ok);
            CHECK(out.ok);
            CHECK_EQ(out.world, static_cast<uint32_t>(world));
            CHECK_EQ(static_cast<int>(out.stats.plane), static_cast<int>(want_plane));
            if (in_place) outputs[r] = work;
        });
    for (auto &t : ts) t.join();
    for (int r = 0; r < world; ++r)
        for (size_t i = 0; i < count; ++i) {
            if constexpr (std::is_floating_point_v<T>)
                CHECK(std::fabs(static_cast<double>(outputs[r][i]) - static_cast<double>(expect[i])) < 1e-4);
            else
                CHECK(outputs[r][i] == expect[i]);
        }
    for (int r = 1; r < world; ++r) CHECK(std::memcmp(outputs[r].data(), outputs[0].data(), count * sizeof(T)) == 0); // bit-identical everywhere
}
} // namespace

TEST(two_level_ring_two_hosts_of_two) {
    TestMaster tm;
    auto peers = connect_on_hosts(tm, {11, 22, 11, 22});
    for (auto &p : peers) CHECK(p->hierarchy_shape() == std::make_pair(2u, 2u));
    reduce_everywhere<float>(peers, 100003, wire::RedOp::Sum, false, pcclExtPlaneTcpHierarchical);
    reduce_everywhere<float>(peers, 100003, wire::RedOp::Avg, true, pcclExtPlaneTcpHierarchical);
    reduce_everywhere<int32_t>(peers, 1025, wire::RedOp::Max, false, pcclExtPlaneTcpHierarchical);
    reduce_everywhere<int64_t>(peers, 3, wire::RedOp::Avg, true, pcclExtPlaneTcpHierarchical); // fewer elements than peers: empty chunks
    reduce_everywhere<double>(peers, 1, wire::RedOp::Min, false, pcclExtPlaneTcpHierarchical);
}

TEST(two_level_ring_three_hosts_of_two_and_two_hosts_of_three) {
    {
        TestMaster tm;
        auto peers = connect_on_hosts(tm, {1, 2, 3, 3, 2, 1});
        for (auto &p : peers) CHECK(p->hierarchy_shape() == std::make_pair(3u, 2u));
        reduce_everywhere<float>(peers, 70001, wire::RedOp::Avg, true, pcclExtPlaneTcpHierarchical);
        reduce_everywhere<uint8_t>(peers, 5000, wire::RedOp::Sum, false, pcclExtPlaneTcpHierarchical);
    }
    {
        TestMaster tm;
        auto peers = connect_on_hosts(tm, {7, 7, 8, 8, 7, 8});
        for (auto &p : peers) CHECK(p->hierarchy_shape() == std::make_pair(2u, 3u));
        reduce_everywhere<double>(peers, 33333, wire::RedOp::Sum, false, pcclExtPlaneTcpHierarchical);
        reduce_everywhere<int16_t>(peers, 7, wire::RedOp::Max, true, pcclExtPlaneTcpHierarchical);
    }
}

TEST(uneven_hosts_keep_the_flat_ring) {
    TestMaster tm;
    auto peers = connect_on_hosts(tm, {1, 1, 1, 2});
    for (auto &p : peers) CHECK(p->hierarchy_shape() == std::make_pair(0u, 0u));
    reduce_everywhere<float>(peers, 4097, wire::RedOp::Sum, false, pcclExtPlaneTcpRing);
}

// Quantised operations on a two-level ring: plain values inside a host, codes between hosts.  Every peer ends with the same
// bits; the traffic of a peer is the local exchange in full precision plus the quantised exchange of its 1/L share.
TEST(two_level_ring_quantises_between_hosts_only) {
    TestMaster tm;
    auto peers = connect_on_hosts(tm, {1, 2, 1, 2, 2, 1}); // 2 hosts x 3 peers
    for (auto &p : peers) CHECK(p->hierarchy_shape() == std::make_pair(2u, 3u));
    const int world = 6;
    const size_t n = 600'011;
    std::vector<std::vector<float>> in(world, std::vector<float>(n));
    for (int r = 0; r < world; ++r)
        for (size_t i = 0; i < n; ++i) in[r][i] = std::cos(0.001f * static_cast<float>(i % 50'000) + 0.5f * r) * 2.0f;
    struct Case {
        wire::QuantAlgo algo;
        wire::DType code;
        wire::RedOp op;
        float tol;
    };
    for (const Case c : {Case{wire::QuantAlgo::ZeroPointScale, wire::DType::U8, wire::RedOp::Sum, 0.2f}, Case{wire::QuantAlgo::MinMax, wire::DType::U8, wire::RedOp::Avg, 0.05f},
                         Case{wire::QuantAlgo::BlockFp8, wire::DType::F8E4M3, wire::RedOp::Avg, 0.2f}}) {
        std::vector<std::vector<float>> out(world, std::vector<float>(n));
        std::vector<uint64_t> tx(world);
        std::vector<std::thread> ts;
        for (int r = 0; r < world; ++r)
            ts.emplace_back([&, r] {
                ReduceRequest req;
                req.send = in[r].data();
                req.recv = out[r].data();
                req.count = n;
                req.op = c.op;
                req.quant = c.algo;
                req.quant_dtype = c.code;
                req.tag = 6;
                ReduceOutcome o;
                CHECK(peers[r]->all_reduce_sync(req, o) == client::Client::Launch::Ok);
                CHECK(o.ok);
                CHECK_EQ(static_cast<int>(o.stats.plane), static_cast<int>(pcclExtPlaneTcpHierarchical));
                tx[r] = o.tx_bytes;
            });
        for (auto &t : ts) t.join();
        for (int r = 1; r < world; ++r) CHECK(std::memcmp(out[0].data(), out[r].data(), n * sizeof(float)) == 0);
        for (size_t i = 0; i < n; i += 7) {
            float truth = 0;
            for (int r = 0; r < world; ++r) truth += in[r][i];
            if (c.op == wire::RedOp::Avg) truth /= world;
            CHECK(std::fabs(out[0][i] - truth) < c.tol);
        }
        // local ring of 3: 2 x 2/3 of the tensor in fp32; cross ring of 2: 2 x 1/2 of the owned third in 8-bit codes
        const double local_bytes = 2.0 * (2.0 / 3.0) * n * 4, cross_bytes = 2.0 * 0.5 * (n / 3.0) * 1.07;
        for (int r = 0; r < world; ++r) CHECK(std::fabs(static_cast<double>(tx[r]) - (local_bytes + cross_bytes)) < 0.05 * (local_bytes + cross_bytes));
    }
    // the constants of the reference's anchor test come out exact through both levels: 6 peers x 42 = 252
    std::vector<std::vector<float>> out(world, std::vector<float>(2048));
    std::vector<std::thread> ts;
    for (int r = 0; r < world; ++r)
        ts.emplace_back([&, r] {
            std::vector<float> x(2048, 42.0f);
            ReduceRequest req;
            req.send = x.data(), req.recv = out[r].data(), req.count = x.size(), req.op = wire::RedOp::Sum;
            req.quant = wire::QuantAlgo::ZeroPointScale, req.quant_dtype = wire::DType::U8, req.tag = 7;
            ReduceOutcome o;
            CHECK(peers[r]->all_reduce_sync(req, o) == client::Client::Launch::Ok && o.ok);
        });
    for (auto &t : ts) t.join();
    for (int r = 0; r < world; ++r)
        for (float v : out[r]) CHECK(v == 252.0f);
}

// ---- lease extension of the collective vote, spoken by hand: two scripted peers on raw sockets (a real peer only offers a
// lease when it has a device plane, which this CPU test does not) -----------------------------------------------------
namespace {
ccoip_socket_address_t loopback(uint16_t port) {
    ccoip_socket_address_t a{};
    a.inet.protocol = inetIPv4;
    a.inet.ipv4.data[0] = 127;
    a.inet.ipv4.data[3] = 1;
    a.port = port;
    return a;
}
struct ScriptedPeer {
    net::BlockingSocket sock;
    Uuid uuid;
    explicit ScriptedPeer(uint16_t master_port) : sock(loopback(master_port)) {
        CHECK(sock.connect(5000));
        wire::C2MSessionRegistration reg;
        reg.p2p_port = 1;
        reg.ss_port = 2;
        reg.bm_port = 3;
        CHECK(sock.send_packet(reg));
        const auto resp = sock.recv_packet<wire::M2CSessionRegistrationResponse>(5000);
        CHECK(resp && resp->accepted);
        uuid = resp->assigned_uuid;
    }
    void finish_round() { // answer the connection info with "links established", take the verdict
        CHECK(sock.recv_packet<wire::M2CP2PConnectionInfo>(5000).has_value());
        wire::C2MP2PEstablished ok;
        ok.success = true;
        CHECK(sock.send_packet(ok));
        const auto verdict = sock.recv_packet<wire::M2CP2PEstablished>(5000);
        CHECK(verdict && verdict->success);
    }
    void vote(uint64_t tag, uint8_t bits) {
        wire::C2MCollectiveInitiate v;
        v.tag = tag;
        v.count = 4;
        v.op = static_cast<wire::RedOp>(static_cast<uint8_t>(wire::RedOp::Sum) | bits);
        CHECK(sock.send_packet(v));
    }
    uint64_t commence(uint64_t tag) {
        const auto go = sock.recv_packet<wire::M2CCollectiveCommence>(5000);
        CHECK(go && go->tag == tag);
        return go->seq_nr;
    }
    void complete(uint64_t tag) {
        wire::C2MCollectiveComplete done;
        done.tag = tag;
        CHECK(sock.send_packet(done));
    }
    void finish(uint64_t tag) {
        const auto ab = sock.recv_packet<wire::M2CCollectiveAbort>(5000);
        CHECK(ab && ab->tag == tag && !ab->aborted);
        const auto done = sock.recv_packet<wire::M2CCollectiveComplete>(5000);
        CHECK(done && done->tag == tag);
    }
};
} // namespace

TEST(master_grants_and_revokes_leases) {
    TestMaster tm;
    ScriptedPeer a(tm.port);
    a.finish_round(); // the first peer is admitted at once
    std::unique_ptr<ScriptedPeer> b;
    std::thread joiner([&] {
        b = std::make_unique<ScriptedPeer>(tm.port);
        b->finish_round();
    });
    std::this_thread::sleep_for(std::chrono::milliseconds(200));
    wire::C2MEstablishP2P admit;
    admit.accept_new_peers = true;
    CHECK(a.sock.send_packet(admit));
    a.finish_round();
    joiner.join();

    auto run_op = [&](uint64_t tag, uint8_t bits_a, uint8_t bits_b) {
        a.vote(tag, bits_a);
        b->vote(tag, bits_b);
        const uint64_t sa = a.commence(tag), sb = b->commence(tag);
        CHECK_EQ(sa, sb);
        a.complete(tag);
        b->complete(tag);
        a.finish(tag);
        b->finish(tag);
        return sa;
    };
    // device plane + lease offer + multicast view from both: granted (bits 63, 61, 60; NVLS bit 62 was not offered)
    uint64_t s = run_op(7, 0x80 | 0x20 | 0x10, 0x80 | 0x20 | 0x10);
    CHECK((s >> 63) == 1 && ((s >> 62) & 1) == 0 && ((s >> 61) & 1) == 1 && ((s >> 60) & 1) == 1);
    // one peer without the multicast view: lease yes, multicast no
    s = run_op(8, 0x80 | 0x20 | 0x10, 0x80 | 0x20);
    CHECK(((s >> 61) & 1) == 1 && ((s >> 60) & 1) == 0);
    // one peer does not offer a lease (or cannot use the device plane): no lease, and a reference peer never sets any of the bits
    s = run_op(9, 0x80 | 0x20, 0x80);
    CHECK((s >> 63) == 1 && ((s >> 61) & 1) == 0);
    s = run_op(10, 0x80 | 0x20, 0);
    CHECK((s >> 60) == 0);

    // a vote on a leased tag ends that tag's lease for everybody -- and only that tag's -- before the vote proceeds
    a.vote(7, 0);
    {
        const auto rv = a.sock.recv_packet<wire::M2CLeaseRevoked>(5000);
        CHECK(rv && rv->tag == 7 && !rv->all && !rv->peer_lost);
        const auto rv2 = b->sock.recv_packet<wire::M2CLeaseRevoked>(5000);
        CHECK(rv2 && rv2->tag == 7 && !rv2->all);
    }
    b->vote(7, 0);
    s = a.commence(7);
    CHECK_EQ(b->commence(7), s);
    CHECK((s >> 60) == 0); // host buffers this time: no device plane, no new lease
    a.complete(7);
    b->complete(7);
    a.finish(7);
    b->finish(7);
    // voting again on the now unleased tag revokes nothing
    s = run_op(7, 0x80 | 0x20, 0x80 | 0x20);
    CHECK(((s >> 61) & 1) == 1);

    // a member vanishes: the survivor is told that every lease is gone and why
    b.reset();
    const auto lost = a.sock.recv_packet<wire::M2CLeaseRevoked>(5000);
    CHECK(lost && lost->all && lost->peer_lost);
}

TEST_MAIN()
