// Asymmetric travelling-salesman solver used to order the ring.  Replaces the reference's libtsp
// (reference: third_party/libtsp/include/libtsp.h:221-243, src/held_karp_tsp.cpp, src/heuristic_tsp.cpp)
// with a from-scratch implementation: exact Held-Karp dynamic programme for small instances and a
// multi-start nearest-neighbour + Or-opt / 2-opt(segment reversal with asymmetric re-costing) /
// 3-opt(segment insertion) local search with a wall-clock budget for larger ones.
#pragma once
#include <cstdint>
#include <limits>
#include <vector>

namespace pccl::topo {

constexpr double kNoEdge = std::numeric_limits<double>::infinity();

struct AtspResult {
    std::vector<int> tour; // permutation of 0..n-1, tour[0] == 0
    double cost = kNoEdge; // sum of edge costs including the closing edge
    bool optimal = false;  // proven optimal (exact solver ran)
    bool feasible = false; // no missing edge used
};

// cost is an n*n row-major matrix, cost[i*n+j] = price of i -> j, kNoEdge when unusable.
AtspResult atsp_solve_exact(const std::vector<double> &cost, int n);
AtspResult atsp_solve_heuristic(const std::vector<double> &cost, int n, int time_limit_ms, int restarts, uint64_t seed,
                                const std::vector<int> *initial_tour = nullptr);
double atsp_tour_cost(const std::vector<double> &cost, int n, const std::vector<int> &tour);

} // namespace pccl::topo
