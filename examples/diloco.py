#!/usr/bin/env python3
"""DiLoCo training of a GPT-2 style model on synthetic tokens (there is no dataset in this sandbox).

Counterpart of the reference's example scripts python/examples/nanogpt_diloco/{sync_diloco,sync_diloco_concat,
async_diloco}.py: every peer trains H inner AdamW steps, then the peers average their pseudo-gradients and
apply an outer Nesterov step.  Differences: parameters and the whole outer step stay on the GPU; peers on one
NVLink domain reduce with the fused kernels.

  # terminal 1 (or pass --spawn-master on rank 0)
  python -m pccl_b200.master --listen-address 0.0.0.0:48148
  # one process per GPU / host
  python examples/diloco.py --master 127.0.0.1:48148 --device cuda:0 [--mode async] [--model tiny]
Peers may be started and killed at any time.
"""
import argparse
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import pccl_b200 as pccl  # noqa: E402
from pccl_b200.models import GPT, GPTConfig, flatten_parameters  # noqa: E402
from pccl_b200.parallel import AsyncDiLoCo, DiLoCo  # noqa: E402
from pccl_b200.utils import Profiler  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--master", default="127.0.0.1:48148")
    ap.add_argument("--spawn-master", action="store_true")
    ap.add_argument("--device", default="cuda:0" if torch.cuda.is_available() else "cpu")
    ap.add_argument("--model", default="tiny", choices=["tiny", "gpt2"])
    ap.add_argument("--mode", default="sync", choices=["sync", "async"])
    ap.add_argument("--inner-steps", type=int, default=8)
    ap.add_argument("--outer-steps", type=int, default=5)
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--min-peers", type=int, default=1)
    ap.add_argument("--quant", default="none", choices=["none", "u8", "fp8"])
    args = ap.parse_args()

    dev = torch.device(args.device)
    if dev.type == "cuda":
        torch.cuda.set_device(dev)
    master = None
    if args.spawn_master:
        master = pccl.MasterNode(args.master)
        master.run()
    comm = pccl.Communicator(args.master)
    comm.connect(n_attempts=30)
    while comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < args.min_peers:
        comm.update_topology()

    torch.manual_seed(0)  # identical initial weights everywhere; late joiners get them via shared state
    cfg = GPTConfig.tiny() if args.model == "tiny" else GPTConfig.gpt2_124m()
    model = GPT(cfg).to(dev)
    inner_flat = flatten_parameters(model)           # the model trains on views into this buffer
    outer = inner_flat.clone()                       # outer ("global") parameters
    quant = {"none": None, "u8": pccl.QuantizationOptions(pccl.DataType.UINT8, pccl.QuantizationAlgorithm.MIN_MAX),
             "fp8": pccl.QuantizationOptions(pccl.DataType.FLOAT8_E4M3, pccl.QuantizationAlgorithm.BLOCK_FP8)}[args.quant]
    algo = (AsyncDiLoCo if args.mode == "async" else DiLoCo)(comm, outer, lr=0.7, momentum=0.9, quantization=quant)
    opt = torch.optim.AdamW(model.parameters(), lr=3e-4)
    step_t = torch.zeros(1, dtype=torch.int64)
    state = pccl.SharedState([pccl.TensorInfo.from_torch(outer, "outer_params"), pccl.TensorInfo.from_torch(algo.buf, "outer_momentum"),
                              pccl.TensorInfo.from_torch(step_t, "outer_step")])
    prof = Profiler(sync_cuda=dev.type == "cuda")
    gen = torch.Generator().manual_seed(1234 + int(time.time() * 1000) % 1000)

    for outer_step in range(args.outer_steps):
        with prof.section("membership"):
            if comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) > 1:
                if comm.are_peers_pending():
                    if args.mode == "async":
                        algo.flush()
                    comm.update_topology()
            else:
                comm.update_topology()
        with prof.section("shared_state"):
            if comm.get_attribute(pccl.Attribute.PEER_GROUP_WORLD_SIZE) > 1:
                if args.mode == "async":
                    algo.flush()
                info = comm.sync_shared_state(state)
                if info.rx_bytes:
                    inner_flat.copy_(outer)          # joined late: start from the group's parameters
                state.push_revision()
        with prof.section("inner"):
            for _ in range(args.inner_steps):
                x = torch.randint(0, cfg.vocab_size, (args.batch, cfg.block_size), generator=gen).to(dev)
                _, loss = model(x, x)
                opt.zero_grad(set_to_none=True)
                loss.backward()
                opt.step()
        with prof.section("outer"):
            algo.outer_step(inner_flat)
            step_t += 1
        world = comm.get_attribute(pccl.Attribute.PEER_GROUP_WORLD_SIZE)
        plane = comm.ext_last_reduce_stats(0)["plane"] if world > 1 and args.mode == "sync" else "-"
        print(f"outer step {outer_step}: loss {loss.item():.4f} world {world} plane {plane}", flush=True)
    if args.mode == "async":
        algo.flush()
    print(prof.report())
    comm.destroy()
    if master:
        master.interrupt()


if __name__ == "__main__":
    main()
