#include "bandwidth_store.hpp"

#include <algorithm>
#include <iomanip>
#include <sstream>

namespace pccl::topo {

bool BandwidthStore::register_peer(const Uuid &peer) { return peers_.insert(peer).second; }

bool BandwidthStore::unregister_peer(const Uuid &peer) {
    if (!peers_.erase(peer)) return false;
    mbps_.erase(peer);
    for (auto &[from, row] : mbps_) row.erase(peer);
    return true;
}

bool BandwidthStore::store(const Uuid &from, const Uuid &to, double mbits_per_second) {
    if (from == to || !peers_.count(from) || !peers_.count(to)) return false;
    mbps_[from][to] = mbits_per_second;
    return true;
}

std::optional<double> BandwidthStore::get(const Uuid &from, const Uuid &to) const {
    const auto row = mbps_.find(from);
    if (row == mbps_.end()) return std::nullopt;
    const auto cell = row->second.find(to);
    if (cell == row->second.end()) return std::nullopt;
    return cell->second;
}

std::vector<Edge> BandwidthStore::missing_edges(const Uuid &peer) const {
    std::vector<Edge> out;
    if (!peers_.count(peer)) return out;
    for (const auto &other : peers_) {
        if (other == peer) continue;
        if (!get(peer, other)) out.push_back({peer, other});
        if (!get(other, peer)) out.push_back({other, peer});
    }
    return out;
}

bool BandwidthStore::fully_populated() const {
    for (const auto &a : peers_)
        for (const auto &b : peers_)
            if (a != b && !get(a, b)) return false;
    return true;
}

std::string BandwidthStore::debug_matrix() const {
    std::vector<Uuid> order(peers_.begin(), peers_.end());
    std::sort(order.begin(), order.end());
    std::ostringstream os;
    os << "bandwidth matrix [Gbit/s], row = sender\n";
    for (const auto &a : order) {
        os << a.str().substr(0, 8) << " |";
        for (const auto &b : order) {
            const auto v = a == b ? std::optional<double>{} : get(a, b);
            if (v)
                os << std::setw(9) << std::fixed << std::setprecision(2) << *v / 1000.0;
            else
                os << std::setw(9) << (a == b ? "-" : "?");
        }
        os << "\n";
    }
    return os.str();
}

} // namespace pccl::topo
