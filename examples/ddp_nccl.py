#!/usr/bin/env python3
"""The baseline the reference ships next to its own DDP example: plain `DistributedDataParallel` over NCCL (gloo on
CPUs), no fault tolerance (reference: python/examples/nanogptddp/train_nccl.py).  Same model, same synthetic tokens and
the same per-step printout as examples/ddp.py, so the two can be compared line by line.

  torchrun --nproc-per-node 8 examples/ddp_nccl.py --model gpt2
"""
import argparse
import os
import sys
import time
from pathlib import Path

import torch
import torch.distributed as dist
from torch.nn.parallel import DistributedDataParallel

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pccl_b200.models import GPT, GPTConfig  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--device", default="cuda" if torch.cuda.is_available() else "cpu")
    ap.add_argument("--model", default="tiny", choices=["tiny", "gpt2"])
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--batch", type=int, default=4)
    args = ap.parse_args()
    on_gpu = args.device.startswith("cuda")
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local_rank) if on_gpu else torch.device("cpu")
    if on_gpu:
        torch.cuda.set_device(dev)
    dist.init_process_group("nccl" if on_gpu else "gloo")
    torch.manual_seed(0)
    cfg = GPTConfig.tiny() if args.model == "tiny" else GPTConfig.gpt2_124m()
    model = DistributedDataParallel(GPT(cfg).to(dev), device_ids=[local_rank] if on_gpu else None)
    opt = torch.optim.AdamW(model.parameters(), lr=3e-4)
    gen = torch.Generator().manual_seed(1234 + dist.get_rank())
    for step in range(args.steps):
        t0 = time.perf_counter()
        x = torch.randint(0, cfg.vocab_size, (args.batch, cfg.block_size), generator=gen).to(dev)
        _, loss = model(x, x)
        opt.zero_grad(set_to_none=True)
        loss.backward()  # bucketed all-reduce of the gradients happens in here
        opt.step()
        if on_gpu:
            torch.cuda.synchronize()
        if dist.get_rank() == 0:
            print(f"step {step}: loss {loss.item():.4f} world {dist.get_world_size()} {1e3 * (time.perf_counter() - t0):.1f} ms", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
