#!/usr/bin/env python3
"""Data-parallel training with gradient averaging through the communicator (reference:
python/examples/nanogptddp/train_pccl.py).  Synthetic tokens, GPT-2 style model.

  python examples/ddp.py --master 127.0.0.1:48148 --spawn-master --device cuda:0 --min-peers 1
"""
import argparse
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import pccl_b200 as pccl  # noqa: E402
from pccl_b200.models import GPT, GPTConfig  # noqa: E402
from pccl_b200.parallel import GradientAverager  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--master", default="127.0.0.1:48148")
    ap.add_argument("--spawn-master", action="store_true")
    ap.add_argument("--device", default="cuda:0" if torch.cuda.is_available() else "cpu")
    ap.add_argument("--model", default="tiny", choices=["tiny", "gpt2"])
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--min-peers", type=int, default=1)
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
    torch.manual_seed(0)
    cfg = GPTConfig.tiny() if args.model == "tiny" else GPTConfig.gpt2_124m()
    model = GPT(cfg).to(dev)
    opt = torch.optim.AdamW(model.parameters(), lr=3e-4)
    avg = GradientAverager(comm, model.parameters(), symmetric=dev.type == "cuda")
    state = pccl.SharedState([pccl.TensorInfo.from_torch(p.data, n) for n, p in model.named_parameters()])
    for step in range(args.steps):
        if comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) > 1 and comm.are_peers_pending():
            comm.update_topology()
        if comm.get_attribute(pccl.Attribute.PEER_GROUP_WORLD_SIZE) > 1:
            comm.sync_shared_state(state)
            state.push_revision()
        x = torch.randint(0, cfg.vocab_size, (args.batch, cfg.block_size)).to(dev)
        _, loss = model(x, x)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        shared = avg.average()
        opt.step()
        print(f"step {step}: loss {loss.item():.4f} averaged={shared}", flush=True)
    comm.destroy()
    if master:
        master.interrupt()


if __name__ == "__main__":
    main()
