#!/usr/bin/env python3
"""One process of the wire-compatibility test: a master or a peer of either implementation.

  interop_peer.py master <ours|ref> <port>
  interop_peer.py peer   <ours|ref> <port> <rank> <world>

`ref` imports the unmodified reference package from baseline/_ref, `ours` this repository's; the Python surface
is the same, so the body below is shared.  A peer prints "PEER OK" when every check passed."""
import sys
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
role, impl, port = sys.argv[1], sys.argv[2], int(sys.argv[3])
if impl == "ref":
    sys.path.insert(0, str(REPO / "baseline" / "_ref"))
    import pccl
else:
    sys.path.insert(0, str(REPO))
    import pccl_b200 as pccl
import numpy as np  # noqa: E402
import torch  # noqa: E402

if role == "master":
    m = pccl.MasterNode(f"0.0.0.0:{port}")
    m.run()
    print("MASTER UP", flush=True)
    sys.stdin.readline()  # parent closes stdin (or writes a line) to stop
    m.interrupt()
    sys.exit(0)

rank, world = int(sys.argv[4]), int(sys.argv[5])
base = 48700 + 8 * rank
comm = pccl.Communicator(f"127.0.0.1:{port}", 0, p2p_connection_pool_size=2, p2p_listen_port=base, shared_state_listen_port=base + 1,
                         benchmark_listen_port=base + 2)
comm.connect(n_attempts=30)
while comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < world:
    comm.update_topology()
    time.sleep(0.01)

# 1. plain reductions, several types and operators
for k, (dtype, op, want) in enumerate([
        (torch.float32, pccl.ReduceOp.SUM, lambda i: sum(r + 1 for r in range(world)) * (i % 7)),
        (torch.int32, pccl.ReduceOp.MAX, lambda i: world * (i % 5)),
        (torch.float64, pccl.ReduceOp.AVG, lambda i: sum(r + 1 for r in range(world)) * (i % 7) / world)]):
    n = 100_003
    idx = torch.arange(n)
    x = ((rank + 1) * (idx % (5 if dtype == torch.int32 else 7))).to(dtype)
    y = torch.empty_like(x)
    info = comm.all_reduce(x, y, op=op, tag=k)
    assert info.local_world_size == world
    expect = torch.tensor([want(i) for i in range(50)], dtype=torch.float64)
    assert torch.allclose(y[:50].double(), expect), (k, y[:10], expect[:10])

# 2. quantised reduction (uint8 min-max): every peer must end with bit-identical output
n = 4096
x = torch.linspace(-1, 1, n) * (rank + 1)
y = torch.empty_like(x)
q = pccl.QuantizationOptions(pccl.DataType.UINT8, pccl.QuantizationAlgorithm.MIN_MAX)
comm.all_reduce(x, y, op=pccl.ReduceOp.SUM, tag=9, quantization_options=q)
truth = torch.linspace(-1, 1, n) * sum(r + 1 for r in range(world))
assert (y - truth).abs().max().item() < 0.05 * world, (y - truth).abs().max().item()
print("QHASH", int(np.frombuffer(y.numpy().tobytes(), dtype=np.uint32).astype(np.uint64).sum()), flush=True)

# 3. shared state: rank 0 sends, the others receive (strategies make the outcome independent of vote arrival order,
#    which the reference master's popularity election is not for a two-way tie)
w = torch.full((50_000,), 3.25 if rank == 0 else -1.0)
state = pccl.SharedState([pccl.TensorInfo.from_torch(w, "w", allow_content_inequality=False)])
state.revision = 1
info = comm.sync_shared_state(state, pccl.SharedStateSyncStrategy.SEND_ONLY if rank == 0 else pccl.SharedStateSyncStrategy.RECEIVE_ONLY)
assert float(w[0]) == 3.25 and float(w[-1]) == 3.25, w[:4]
assert (info.rx_bytes == 0) == (rank == 0), (rank, info.tx_bytes, info.rx_bytes)
# 4. everybody identical now: the popularity election finds one candidate and nothing moves
state.revision = 2
info = comm.sync_shared_state(state, pccl.SharedStateSyncStrategy.ENFORCE_POPULAR)
assert info.rx_bytes == 0 and info.tx_bytes == 0, (info.tx_bytes, info.rx_bytes)
# 4b. (optional) a state large enough for this implementation to fetch it over several connections at once -- from a
#     distributor of either implementation, which sees nothing but independent plain requests
if len(sys.argv) > 6 and sys.argv[6] == "bigstate":
    tensors = [torch.full((1_500_000 + 1000 * k,), float(k) if rank == 0 else -7.0) for k in range(12)]  # 72 MB
    big = pccl.SharedState([pccl.TensorInfo.from_torch(t, f"layer{k}") for k, t in enumerate(tensors)])
    big.revision = 3
    info = comm.sync_shared_state(big, pccl.SharedStateSyncStrategy.SEND_ONLY if rank == 0 else pccl.SharedStateSyncStrategy.RECEIVE_ONLY)
    for k, t in enumerate(tensors):
        assert float(t[0]) == float(k) and float(t[-1]) == float(k) and float(t[len(t) // 2]) == float(k), (k, t[:3])
    total = sum(t.numel() * 4 for t in tensors)
    assert info.rx_bytes == (0 if rank == 0 else total), (rank, info.rx_bytes, total)
# 5. (optional) topology optimisation: every peer measures its bandwidth to the others with the benchmark protocol
if len(sys.argv) > 6 and sys.argv[6] == "topology":
    comm.optimize_topology()
    x = torch.ones(1000)
    comm.all_reduce(x, x, op=pccl.ReduceOp.SUM, tag=77)
    assert float(x[0]) == world
print("PEER OK", flush=True)
