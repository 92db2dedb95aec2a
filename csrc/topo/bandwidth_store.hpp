// Directed bandwidth matrix between the peers of one peer group, filled from benchmark reports
// (reference: ccoip/internal/bandwidth_store.hpp:16-104).
#pragma once
#include <optional>
#include <string>
#include <unordered_map>
#include <unordered_set>
#include <vector>

#include "../util/uuid.hpp"

namespace pccl::topo {

struct Edge {
    Uuid from, to;
    bool operator==(const Edge &o) const { return from == o.from && to == o.to; }
};

class BandwidthStore {
public:
    bool register_peer(const Uuid &peer);
    bool unregister_peer(const Uuid &peer); // also forgets every edge touching it
    bool store(const Uuid &from, const Uuid &to, double mbits_per_second);
    std::optional<double> get(const Uuid &from, const Uuid &to) const;
    // all unmeasured edges that start or end at `peer`
    std::vector<Edge> missing_edges(const Uuid &peer) const;
    bool fully_populated() const;
    size_t num_peers() const { return peers_.size(); }
    const std::unordered_set<Uuid> &peers() const { return peers_; }
    std::string debug_matrix() const; // Gbit/s table

private:
    std::unordered_set<Uuid> peers_;
    std::unordered_map<Uuid, std::unordered_map<Uuid, double>> mbps_;
};

} // namespace pccl::topo
