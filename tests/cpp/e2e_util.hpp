// In-process cluster helper: one master on an ephemeral port plus N clients on threads.
#pragma once
#include <chrono>
#include <memory>
#include <thread>
#include <vector>

#include "client/client.hpp"
#include "harness.hpp"
#include "master/master.hpp"
#include "net/addr.hpp"

struct TestMaster {
    std::unique_ptr<pccl::master::Master> m;
    uint16_t port = 0;
    TestMaster() {
        ccoip_socket_address_t a{};
        a.inet.protocol = inetIPv4;
        a.inet.ipv4.data[0] = 127;
        a.inet.ipv4.data[3] = 1;
        a.port = 0;
        m = std::make_unique<pccl::master::Master>(a);
        CHECK(m->run());
        port = m->port();
    }
    ~TestMaster() {
        m->interrupt();
        m->join();
    }
    uint32_t pool_size = 1;
    pccl::client::ClientConfig config(uint32_t group = 0) const {
        pccl::client::ClientConfig c;
        c.master.inet.protocol = inetIPv4;
        c.master.inet.ipv4.data[0] = 127;
        c.master.inet.ipv4.data[3] = 1;
        c.master.port = port;
        c.peer_group = group;
        c.pool_size = pool_size;
        return c;
    }
};

// Connects `n` clients.  Already connected peers keep calling update_topology() until the
// newcomer is in, the way an application's training loop would.
// `hosts` (one id per client) turns the two-level ring on and pretends the clients run on that many machines.
inline std::vector<std::unique_ptr<pccl::client::Client>> connect_peers(const TestMaster &tm, int n, const std::vector<uint32_t> &groups = {},
                                                                        const std::vector<uint64_t> &hosts = {}) {
    std::vector<std::unique_ptr<pccl::client::Client>> peers;
    for (int i = 0; i < n; ++i) {
        auto cfg = tm.config(groups.empty() ? 0 : groups[i]);
        if (!hosts.empty()) cfg.host_id = hosts[i], cfg.hierarchy = 1;
        auto c = std::make_unique<pccl::client::Client>(cfg);
        std::atomic<bool> joined{false};
        std::thread joiner([&] {
            CHECK(c->connect());
            joined = true;
        });
        while (!joined) {
            if (!peers.empty() && peers[0]->global_world() < static_cast<uint64_t>(i + 1)) {
                std::vector<std::thread> ts;
                for (auto &p : peers) ts.emplace_back([&p] { CHECK(p->update_topology()); });
                for (auto &t : ts) t.join();
            } else {
                std::this_thread::sleep_for(std::chrono::milliseconds(2));
            }
        }
        joiner.join();
        peers.push_back(std::move(c));
    }
    return peers;
}
