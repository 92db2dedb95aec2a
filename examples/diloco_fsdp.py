#!/usr/bin/env python3
"""DiLoCo across "islands" whose model is sharded with FSDP2 inside each island.

Counterpart of the reference's python/examples/nanogpt_diloco/sync_diloco_fsdp.py: inside one island (one torchrun
job: a node, or a few GPUs) the model is sharded with `fully_shard` and trained with the island's own process
group (NCCL on GPUs, gloo on CPUs); across islands the ranks that hold the SAME shard form one PCCL peer group
(`peer_group = rank inside the island`, reference :351) and run the DiLoCo outer step on their shard.  Membership
votes are global, collectives and shared state are per peer group, so islands can come and go as a whole.

  python -m pccl_b200.master --listen-address 0.0.0.0:48148
  # island A                                                          # island B (any time later)
  torchrun --nproc-per-node 4 examples/diloco_fsdp.py --pccl-master HOST:48148
Every rank of an island must reach the outer step with the same set of islands, so the loop only proceeds when the
global world is a whole number of islands (reference: docs/md/8_CommonFootguns.md:1-101).
"""
import argparse
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist
from torch.distributed.fsdp import fully_shard

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import pccl_b200 as pccl  # noqa: E402
from pccl_b200 import ops  # noqa: E402
from pccl_b200.models import GPT, GPTConfig  # noqa: E402
from pccl_b200.parallel import DiLoCo  # noqa: E402


def local_shards(model):
    """The tensors this rank actually stores: FSDP2 parameters are DTensors, PCCL wants the local shard."""
    with torch.no_grad():
        return [(p.detach().to_local() if hasattr(p, "to_local") else p.detach()) for p in model.parameters()]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--pccl-master", dest="master", default="127.0.0.1:48148")
    ap.add_argument("--device", default="cuda" if torch.cuda.is_available() else "cpu")
    ap.add_argument("--model", default="tiny", choices=["tiny", "gpt2"])
    ap.add_argument("--inner-steps", type=int, default=4)
    ap.add_argument("--outer-steps", type=int, default=3)
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--min-islands", type=int, default=1)
    ap.add_argument("--p2p-port-base", type=int, default=48200)
    args = ap.parse_args()

    island_rank, island_size = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    on_gpu = args.device.startswith("cuda")
    dev = torch.device("cuda", local_rank) if on_gpu else torch.device("cpu")
    if on_gpu:
        torch.cuda.set_device(dev)
    dist.init_process_group("nccl" if on_gpu else "gloo")

    torch.manual_seed(0)  # same initial weights on every island
    cfg = GPTConfig.tiny() if args.model == "tiny" else GPTConfig.gpt2_124m()
    model = GPT(cfg).to(dev)
    for block in model.blocks:
        fully_shard(block)
    fully_shard(model)
    opt = torch.optim.AdamW(model.parameters(), lr=3e-4)

    # one communicator per shard rank; the peer group ties equal shards of different islands together
    comm = pccl.Communicator(args.master, peer_group=island_rank, p2p_listen_port=args.p2p_port_base + 8 * island_rank,
                             shared_state_listen_port=args.p2p_port_base + 8 * island_rank + 1,
                             benchmark_listen_port=args.p2p_port_base + 8 * island_rank + 2)
    comm.connect(n_attempts=30)

    def islands_complete() -> bool:
        # every island contributes exactly `island_size` peers: wait until nobody is half-way in
        world = comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE)
        largest = comm.get_attribute(pccl.Attribute.LARGEST_PEER_GROUP_WORLD_SIZE)
        return world == island_size * largest and largest >= args.min_islands

    while not islands_complete():
        comm.update_topology()

    shards = local_shards(model)
    n = sum(s.numel() for s in shards)
    inner = torch.empty(n, dtype=torch.float32, device=dev)   # this rank's shard of the inner parameters, flat
    outer = torch.empty_like(inner)                           # and of the outer parameters

    def gather(dst):
        off = 0
        with torch.no_grad():
            for s in shards:
                dst[off:off + s.numel()].copy_(s.reshape(-1))
                off += s.numel()

    def scatter(src):
        off = 0
        with torch.no_grad():
            for s in shards:
                s.copy_(src[off:off + s.numel()].view_as(s))
                off += s.numel()

    gather(outer)
    algo = DiLoCo(comm, outer, lr=0.7, momentum=0.9)
    step_t = torch.zeros(1, dtype=torch.int64)
    state = pccl.SharedState([pccl.TensorInfo.from_torch(outer, "outer_shard"), pccl.TensorInfo.from_torch(algo.buf, "momentum_shard"),
                              pccl.TensorInfo.from_torch(step_t, "outer_step")])
    gen = torch.Generator().manual_seed(1000 * island_rank + os.getpid() % 1000)

    for outer_step in range(args.outer_steps):
        # membership: admit newcomers, but only continue with whole islands
        if comm.are_peers_pending() or not islands_complete():
            comm.update_topology()
            while not islands_complete():
                comm.update_topology()
        if comm.get_attribute(pccl.Attribute.PEER_GROUP_WORLD_SIZE) > 1:
            info = comm.sync_shared_state(state)
            if info.rx_bytes:
                scatter(outer)  # this island joined late: its shards start from the group's parameters
            state.push_revision()
        for _ in range(args.inner_steps):
            x = torch.randint(0, cfg.vocab_size, (args.batch, cfg.block_size), generator=gen).to(dev)
            _, loss = model(x, x)
            opt.zero_grad(set_to_none=True)
            loss.backward()
            opt.step()
        gather(inner)
        algo.outer_step(inner)  # averages this shard's pseudo-gradient over the islands, Nesterov update
        scatter(inner)          # inner now holds the new outer parameters
        step_t += 1
        digest = ops.simplehash(outer)
        print(f"[island rank {island_rank}] outer step {outer_step}: loss {loss.item():.4f} islands "
              f"{comm.get_attribute(pccl.Attribute.PEER_GROUP_WORLD_SIZE)} shard elems {n} digest {digest}", flush=True)
    dist.barrier()
    comm.destroy()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
