#include "optimizer.hpp"

#include <algorithm>
#include <unordered_map>

#include "../util/log.hpp"
#include "atsp.hpp"

namespace pccl::topo {

namespace {

struct Problem {
    std::vector<Uuid> nodes;
    std::vector<double> cost;
    std::vector<int> current; // current ring as node indices (may be empty if membership differs)
};

Problem build(const BandwidthStore &store, const std::vector<Uuid> &ring) {
    Problem p;
    p.nodes.assign(store.peers().begin(), store.peers().end());
    std::sort(p.nodes.begin(), p.nodes.end());
    const int n = static_cast<int>(p.nodes.size());
    p.cost.assign(static_cast<size_t>(n) * n, kNoEdge);
    for (int i = 0; i < n; ++i)
        for (int j = 0; j < n; ++j) {
            if (i == j) continue;
            if (const auto bw = store.get(p.nodes[i], p.nodes[j]); bw && *bw > 0) p.cost[i * n + j] = 1000.0 / *bw;
        }
    std::unordered_map<Uuid, int> index;
    for (int i = 0; i < n; ++i) index[p.nodes[i]] = i;
    if (static_cast<int>(ring.size()) == n) {
        for (const auto &u : ring) {
            const auto it = index.find(u);
            if (it == index.end()) {
                p.current.clear();
                break;
            }
            p.current.push_back(it->second);
        }
    }
    return p;
}

void emit(const Problem &p, const AtspResult &r, std::vector<Uuid> &ring) {
    ring.clear();
    for (int idx : r.tour) ring.push_back(p.nodes[idx]);
}

} // namespace

bool optimize_ring(const BandwidthStore &store, std::vector<Uuid> &ring, bool &is_optimal) {
    const Problem p = build(store, ring);
    const int n = static_cast<int>(p.nodes.size());
    is_optimal = false;
    if (n == 0) return false;
    AtspResult r = n <= 12 ? atsp_solve_exact(p.cost, n)
                           : atsp_solve_heuristic(p.cost, n, 1000, 4 * n, 0x5eed, p.current.empty() ? nullptr : &p.current);
    if (!r.feasible) return false;
    emit(p, r, ring);
    is_optimal = r.optimal;
    PLOG(INFO) << "ring optimisation: n=" << n << " cost=" << r.cost << (r.optimal ? " (optimal)" : " (heuristic)");
    return true;
}

bool improve_ring_moonshot(const BandwidthStore &store, std::vector<Uuid> &ring, bool &is_optimal, bool &improved) {
    const Problem p = build(store, ring);
    const int n = static_cast<int>(p.nodes.size());
    is_optimal = improved = false;
    if (n == 0) return false;
    const double before = p.current.empty() ? kNoEdge : atsp_tour_cost(p.cost, n, p.current);
    AtspResult r = n <= 16 ? atsp_solve_exact(p.cost, n)
                           : atsp_solve_heuristic(p.cost, n, 30000, 16 * n, 0xa11ce, p.current.empty() ? nullptr : &p.current);
    if (!r.feasible) return false;
    is_optimal = r.optimal;
    if (r.cost + 1e-12 < before) {
        improved = true;
        emit(p, r, ring);
    }
    return true;
}

} // namespace pccl::topo
