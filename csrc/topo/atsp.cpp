#include "atsp.hpp"

#include <algorithm>
#include <chrono>
#include <numeric>
#include <random>

namespace pccl::topo {

double atsp_tour_cost(const std::vector<double> &cost, int n, const std::vector<int> &tour) {
    if (static_cast<int>(tour.size()) != n || n == 0) return kNoEdge;
    if (n == 1) return 0.0;
    double c = 0;
    for (int i = 0; i < n; ++i) c += cost[tour[i] * n + tour[(i + 1) % n]];
    return c;
}

static void rotate_to_zero(std::vector<int> &tour) {
    const auto it = std::find(tour.begin(), tour.end(), 0);
    if (it != tour.end()) std::rotate(tour.begin(), it, tour.end());
}

AtspResult atsp_solve_exact(const std::vector<double> &cost, int n) {
    AtspResult res;
    if (n <= 0) return res;
    if (n == 1) {
        res.tour = {0};
        res.cost = 0;
        res.optimal = res.feasible = true;
        return res;
    }
    // dp[mask][last]: cheapest path that starts at node 0, visits exactly `mask` (node 0 excluded
    // from the mask), and ends in `last`.
    const int m = n - 1;
    const size_t states = size_t(1) << m;
    std::vector<double> dp(states * m, kNoEdge);
    std::vector<int8_t> parent(states * m, -1);
    for (int j = 0; j < m; ++j) dp[(size_t(1) << j) * m + j] = cost[0 * n + (j + 1)];
    for (size_t mask = 1; mask < states; ++mask) {
        for (int last = 0; last < m; ++last) {
            if (!(mask >> last & 1)) continue;
            const double base = dp[mask * m + last];
            if (base == kNoEdge) continue;
            for (int nxt = 0; nxt < m; ++nxt) {
                if (mask >> nxt & 1) continue;
                const double c = base + cost[(last + 1) * n + (nxt + 1)];
                const size_t idx = (mask | size_t(1) << nxt) * m + nxt;
                if (c < dp[idx]) {
                    dp[idx] = c;
                    parent[idx] = static_cast<int8_t>(last);
                }
            }
        }
    }
    const size_t full = states - 1;
    double best = kNoEdge;
    int best_last = -1;
    for (int last = 0; last < m; ++last) {
        const double c = dp[full * m + last] + cost[(last + 1) * n + 0];
        if (c < best) {
            best = c;
            best_last = last;
        }
    }
    if (best_last < 0) return res; // no Hamiltonian cycle over usable edges
    std::vector<int> rev;
    size_t mask = full;
    int cur = best_last;
    while (cur >= 0) {
        rev.push_back(cur + 1);
        const int p = parent[mask * m + cur];
        mask &= ~(size_t(1) << cur);
        cur = p;
    }
    res.tour.push_back(0);
    res.tour.insert(res.tour.end(), rev.rbegin(), rev.rend());
    res.cost = best;
    res.optimal = true;
    res.feasible = true;
    return res;
}

namespace {

struct Search {
    const std::vector<double> &cost;
    int n;
    double c(int a, int b) const { return cost[a * n + b]; }

    // Or-opt: move a segment of length 1..3 to another position (keeps orientation, so it is
    // exact for asymmetric costs).  Returns true when an improving move was applied.
    bool or_opt(std::vector<int> &t, double &total) const {
        for (int len = 1; len <= 3 && len < n - 1; ++len) {
            for (int i = 0; i < n; ++i) {
                // segment t[i..i+len-1] (cyclic)
                const int prev = t[(i - 1 + n) % n], first = t[i], last = t[(i + len - 1) % n], next = t[(i + len) % n];
                const double removed = c(prev, first) + c(last, next) - c(prev, next);
                for (int k = 0; k < n - len - 1; ++k) {
                    const int a = t[(i + len + k) % n], b = t[(i + len + k + 1) % n];
                    const double delta = c(a, first) + c(last, b) - c(a, b) - removed;
                    if (delta < -1e-12) {
                        std::vector<int> seg, rest;
                        for (int s = 0; s < len; ++s) seg.push_back(t[(i + s) % n]);
                        for (int s = 0; s < n - len; ++s) rest.push_back(t[(i + len + s) % n]);
                        std::vector<int> out;
                        for (int s = 0; s <= k; ++s) out.push_back(rest[s]);
                        out.insert(out.end(), seg.begin(), seg.end());
                        for (int s = k + 1; s < n - len; ++s) out.push_back(rest[s]);
                        t = std::move(out);
                        total += delta;
                        return true;
                    }
                }
            }
        }
        return false;
    }

    // 2-opt with full re-costing of the reversed segment (direction matters).
    bool two_opt(std::vector<int> &t, double &total) const {
        for (int i = 0; i < n - 1; ++i) {
            for (int j = i + 1; j < n; ++j) {
                if (j - i >= n - 1) continue;
                const int before = t[(i - 1 + n) % n], after = t[(j + 1) % n];
                double old_c = c(before, t[i]) + c(t[j], after), new_c = c(before, t[j]) + c(t[i], after);
                for (int k = i; k < j; ++k) {
                    old_c += c(t[k], t[k + 1]);
                    new_c += c(t[k + 1], t[k]);
                }
                if (new_c < old_c - 1e-12) {
                    std::reverse(t.begin() + i, t.begin() + j + 1);
                    total += new_c - old_c;
                    return true;
                }
            }
        }
        return false;
    }

    void descend(std::vector<int> &t, double &total, const std::chrono::steady_clock::time_point &deadline) const {
        for (;;) {
            if (std::chrono::steady_clock::now() > deadline) return;
            if (or_opt(t, total)) continue;
            if (two_opt(t, total)) continue;
            return;
        }
    }
};

} // namespace

AtspResult atsp_solve_heuristic(const std::vector<double> &cost, int n, int time_limit_ms, int restarts, uint64_t seed,
                                const std::vector<int> *initial_tour) {
    AtspResult best;
    if (n <= 0) return best;
    if (n <= 2) return atsp_solve_exact(cost, n);
    // Missing edges get a large finite penalty so the search can walk through infeasible tours.
    double max_finite = 0;
    for (double v : cost)
        if (v != kNoEdge) max_finite = std::max(max_finite, v);
    const double penalty = (max_finite + 1.0) * n * 4;
    std::vector<double> work(cost);
    for (auto &v : work)
        if (v == kNoEdge) v = penalty;
    const Search search{work, n};
    std::mt19937_64 rng(seed);
    const auto deadline = std::chrono::steady_clock::now() + std::chrono::milliseconds(time_limit_ms);

    auto consider = [&](std::vector<int> t) {
        double total = atsp_tour_cost(work, n, t);
        search.descend(t, total, deadline);
        total = atsp_tour_cost(work, n, t);
        if (best.tour.empty() || total < best.cost) {
            best.tour = std::move(t);
            best.cost = total;
        }
    };

    if (initial_tour && static_cast<int>(initial_tour->size()) == n) consider(*initial_tour);
    for (int r = 0; r < std::max(1, restarts); ++r) {
        if (r > 0 && std::chrono::steady_clock::now() > deadline) break;
        std::vector<int> t;
        if (r % 2 == 0) {
            // nearest neighbour from a (rotating) start node
            std::vector<char> used(n, 0);
            int cur = (r / 2) % n;
            t.push_back(cur);
            used[cur] = 1;
            for (int step = 1; step < n; ++step) {
                int pick = -1;
                for (int j = 0; j < n; ++j)
                    if (!used[j] && (pick < 0 || work[cur * n + j] < work[cur * n + pick])) pick = j;
                used[pick] = 1;
                t.push_back(pick);
                cur = pick;
            }
        } else {
            t.resize(n);
            std::iota(t.begin(), t.end(), 0);
            std::shuffle(t.begin(), t.end(), rng);
        }
        consider(std::move(t));
    }
    rotate_to_zero(best.tour);
    best.cost = atsp_tour_cost(cost, n, best.tour);
    best.feasible = best.cost != kNoEdge;
    best.optimal = false;
    return best;
}

} // namespace pccl::topo
