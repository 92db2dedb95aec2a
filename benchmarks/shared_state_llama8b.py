#!/usr/bin/env python3
"""Shared-state synchronisation of Llama-3-8B-shaped bf16 weights (BASELINE.json config 4) between peers on
one box, one peer per GPU, under torchrun:

  phase A  rank 0 holds the weights, everybody else joins with garbage -> full transfer, hashes checked
  phase B  nothing changed -> only the GPU hashes run, zero bytes move
  phase C  rank 0 changed ONE tensor -> only that tensor moves
Wall-clock per phase (max over ranks).  `--scale` shrinks every dimension for quick tests.
"""
import argparse
import json
import os
import sys
import time
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import pccl_b200 as pccl  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--scale", type=float, default=1.0)
args = ap.parse_args()
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
box = [0]
master = None
if rank == 0:
    master = pccl.MasterNode("127.0.0.1:0")
    master.run()
    box[0] = master.port
dist.broadcast_object_list(box, src=0)
comm = pccl.Communicator(f"127.0.0.1:{box[0]}")
comm.connect(n_attempts=15)
while comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < world:
    comm.update_topology()

s = args.scale
H, KV, FF, V, L = int(4096 * s), int(1024 * s), int(14336 * s), int(128256 * s), max(1, int(32 * s))
shapes = [("embed_tokens", (V, H)), ("lm_head", (V, H)), ("norm", (H,))]
for i in range(L):
    p = f"layers.{i}."
    shapes += [(p + "q_proj", (H, H)), (p + "k_proj", (KV, H)), (p + "v_proj", (KV, H)), (p + "o_proj", (H, H)),
               (p + "gate_proj", (FF, H)), (p + "up_proj", (FF, H)), (p + "down_proj", (H, FF)),
               (p + "input_layernorm", (H,)), (p + "post_attention_layernorm", (H,))]
torch.manual_seed(0 if rank == 0 else 100 + rank)
tensors = {name: torch.randn(shape, device=dev, dtype=torch.bfloat16) * 0.02 for name, shape in shapes}
n_params = sum(t.numel() for t in tensors.values())
state = pccl.SharedState([pccl.TensorInfo.from_torch(t, name) for name, t in tensors.items()])
state.revision = 1


def phase(strategy):
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    info = comm.sync_shared_state(state, strategy)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()), info


out = {"params": n_params, "bytes": n_params * 2, "tensors": len(tensors), "world": world}
tA, infoA = phase(pccl.SharedStateSyncStrategy.SEND_ONLY if rank == 0 else pccl.SharedStateSyncStrategy.RECEIVE_ONLY)
ref = tensors["layers.0.q_proj"].float().sum().item()
chk = torch.tensor([ref], device=dev, dtype=torch.float64)
dist.broadcast(chk, src=0)
assert abs(chk.item() - ref) < 1e-3, "weights differ after the sync"
state.push_revision()
tB, infoB = phase(pccl.SharedStateSyncStrategy.ENFORCE_POPULAR)
assert infoB.tx_bytes == 0 and infoB.rx_bytes == 0
if rank == 0:
    tensors["layers.0.down_proj"].add_(1.0)
state.push_revision()
tC, infoC = phase(pccl.SharedStateSyncStrategy.SEND_ONLY if rank == 0 else pccl.SharedStateSyncStrategy.RECEIVE_ONLY)
one = tensors["layers.0.down_proj"].numel() * 2
if rank == 1:
    assert infoC.rx_bytes == one, (infoC.rx_bytes, one)
# phase D: the last peer leaves the run and comes back with garbage in its weights (a restarted worker): the others admit it
# and one synchronisation brings it up to date (popular content wins; only the joiner receives)
tD = tD_sync = None
infoD = None
if world >= 2:
    state.push_revision()
    dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    if rank == world - 1:
        comm.destroy()
        for t in tensors.values():
            t.normal_()
        comm = pccl.Communicator(f"127.0.0.1:{box[0]}")
        comm.connect(n_attempts=30)
        state = pccl.SharedState([pccl.TensorInfo.from_torch(t, name) for name, t in tensors.items()])
        state.revision = 0  # knows nothing: whatever the group holds is newer
    else:
        while comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) != world - 1:
            time.sleep(0.001)  # the master has noticed the departure
            comm.update_topology()
        while comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < world:
            comm.update_topology()
    t1 = time.perf_counter()
    infoD = comm.sync_shared_state(state, pccl.SharedStateSyncStrategy.ENFORCE_POPULAR)
    torch.cuda.synchronize()
    t2 = time.perf_counter()
    tt = torch.tensor([t1 - t0, t2 - t1], device=dev, dtype=torch.float64)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    tD, tD_sync = float(tt[0]), float(tt[1])
    ref = tensors["layers.0.q_proj"].float().sum().item()
    chk = torch.tensor([ref], device=dev, dtype=torch.float64)
    dist.broadcast(chk, src=0)
    assert abs(chk.item() - ref) < 1e-3, "the rejoined peer does not hold the group's weights"
    if rank == world - 1:
        assert infoD.rx_bytes == n_params * 2
    out.update({"leave_and_rejoin_membership_s": round(tD, 4), "rejoin_sync_s": round(tD_sync, 4),
                "rejoin_sync_GBps": round(n_params * 2 / tD_sync / 1e9, 1)})
out.update({"full_transfer_s": round(tA, 4), "full_transfer_GBps_per_receiver": round(n_params * 2 / tA / 1e9, 1),
            "rank_rx_bytes_A": infoA.rx_bytes, "rank_tx_bytes_A": infoA.tx_bytes,
            "hash_only_s": round(tB, 4), "hash_only_GBps": round(n_params * 2 / tB / 1e9, 1),
            "one_tensor_s": round(tC, 4), "one_tensor_bytes": one})
if rank == 0:
    print(json.dumps(out), flush=True)
comm.destroy()
dist.barrier()
dist.destroy_process_group()
