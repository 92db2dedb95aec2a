// Ring-order optimisation over a BandwidthStore (reference: ccoip/src/cpp/topolgy_optimizer.cpp:6-182).
// Edge cost is 1000 / Mbit/s (the reference objective: sum of inverse bandwidths along the ring).
#pragma once
#include <vector>

#include "bandwidth_store.hpp"

namespace pccl::topo {

// `ring` carries the current order in and the new order out.  Returns false when no Hamiltonian
// cycle exists over the measured edges.
bool optimize_ring(const BandwidthStore &store, std::vector<Uuid> &ring, bool &is_optimal);
// Long-running improvement pass started from the current ring.
bool improve_ring_moonshot(const BandwidthStore &store, std::vector<Uuid> &ring, bool &is_optimal, bool &improved);

} // namespace pccl::topo
