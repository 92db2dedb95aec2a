#!/usr/bin/env python3
"""Reference arm of bench.py: the UNMODIFIED reference (jundi69/pccl) installed into baseline/_ref,
driven through its own public Python API and stock code path.

Workload = the same as the main arm: fp32 AVG all-reduce of `--params` pseudo-gradients that live on
the GPU.  The reference reduces host memory only (python/framework/pccl/_pccl.py:617), so, exactly like its
own examples (python/examples/nanogpt_diloco/sync_diloco.py:494), each step is D2H into pinned memory ->
`Communicator.all_reduce` -> H2D.  Device time by CUDA events, max over peers.

One peer cannot all-reduce in the reference (pcclTooFewPeers), therefore `--gpus 1` runs its minimum of
two peers, both staging through GPU 0, and says so in `config`.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
REF = REPO / "baseline" / "_ref"
PUBLISHED_BASELINE_GBPS = 45.0 / 8.0


def peer_main(rank: int, world: int, device: int, master_port: int, params: int, steps: int, warmup: int) -> dict:
    sys.path.insert(0, str(REF))
    import pccl  # the reference package
    import torch

    torch.cuda.set_device(device)
    dev = torch.device("cuda", device)
    torch.manual_seed(1234 + rank)
    n = params
    grad = torch.randn(n, device=dev)
    host = torch.empty(n, dtype=torch.float32).pin_memory()
    pinned_in = torch.randn(n).pin_memory()
    base = 48400 + 8 * rank
    comm = pccl.Communicator(f"127.0.0.1:{master_port}", 0, p2p_connection_pool_size=8, p2p_listen_port=base,
                             shared_state_listen_port=base + 1, benchmark_listen_port=base + 2)
    comm.connect(n_attempts=30)
    while comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < world:
        comm.update_topology()

    def step():
        host.copy_(grad, non_blocking=False)                      # D2H
        comm.all_reduce(host, host, op=pccl.ReduceOp.AVG)         # host ring over loopback TCP
        grad.copy_(host, non_blocking=False)                      # H2D

    def e2e_step():
        comm.all_reduce(pinned_in, host, op=pccl.ReduceOp.AVG)    # input already lives in pinned host memory
        grad.copy_(host, non_blocking=False)                      # the result is needed on the device
        return float(grad[0].item())

    def timed(fn, k):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(k):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / k

    for _ in range(max(3, warmup)):
        step()
    ms = timed(step, steps)
    for _ in range(2):
        e2e_step()
    ms_e2e = timed(e2e_step, max(2, steps // 2))
    return {"rank": rank, "ms_per_step": ms, "ms_e2e": ms_e2e}


def run(args) -> dict:
    sys.path.insert(0, str(REF))
    import pccl
    world_env = int(os.environ.get("WORLD_SIZE", "1"))
    n, steps, warmup = args.params, args.steps, args.warmup
    nbytes = n * 4
    if world_env > 1:
        import torch
        import torch.distributed as dist
        rank, local = int(os.environ["RANK"]), int(os.environ["LOCAL_RANK"])
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        port = int(os.environ.get("MASTER_PORT", "29500")) + 53
        master = None
        if rank == 0:
            master = pccl.MasterNode(f"0.0.0.0:{port}")
            master.run()
        dist.barrier()
        r = peer_main(rank, world_env, local, port, n, steps, warmup)
        t = torch.tensor([r["ms_per_step"], r["ms_e2e"]], device=torch.device("cuda", local), dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, ms_e2e = float(t[0].item()), float(t[1].item())
        dist.barrier()
        if master is not None:
            master.interrupt()
        dist.destroy_process_group()
        if rank != 0:
            return {}
        peers, note = world_env, "one reference peer per GPU, tensors staged D2H -> host all-reduce -> H2D"
    else:
        port = 48390
        master = pccl.MasterNode(f"0.0.0.0:{port}")
        master.run()
        procs = [subprocess.Popen([sys.executable, str(Path(__file__).resolve()), "--peer", str(r), "--world", "2", "--device", "0",
                                   "--port", str(port), "--params", str(n), "--steps", str(steps), "--warmup", str(warmup)],
                                  stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True) for r in range(2)]
        outs = [json.loads(p.communicate(timeout=3000)[0].strip().splitlines()[-1]) for p in procs]
        master.interrupt()
        ms = max(o["ms_per_step"] for o in outs)
        ms_e2e = max(o["ms_e2e"] for o in outs)
        peers, note = 2, "the reference needs >= 2 peers: two peers share GPU 0, tensors staged D2H -> host all-reduce -> H2D"
    factor = 2.0 * (peers - 1) / peers
    value = nbytes * factor / (ms * 1e-3) / 1e9
    return {
        "impl": "reference", "metric": "allreduce_bus_GBps", "value": round(value, 4), "unit": "GB/s", "n_gpus": args.gpus,
        "steps": steps, "warmup": max(3, warmup), "ms_per_step": round(ms, 3), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": round(value / PUBLISHED_BASELINE_GBPS, 3), "dtype": "fp32", "data": "synthetic",
        "config": {"model": "DiLoCo pseudo-gradients, GPT-2-class 150M params fp32", "params": n, "parallelism": f"dp{peers}", "op": "avg",
                   "peers": peers, "note": note,
                   "e2e_note": "e2e: the step's input already sits in pinned host memory, so the reference skips the device-to-host half of its "
                               "staging (all-reduce on host memory, then host-to-device copy of the result, then a 4-byte read)", "build": "reference @ c488cabf, PCCL_BUILD_CUDA_SUPPORT=OFF, piquant stand-in header (not exercised)"},
        "e2e": {"value": round(nbytes * factor / (ms_e2e * 1e-3) / 1e9, 4), "unit": "GB/s", "ms_per_step": round(ms_e2e, 3),
                "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": 4},
        "gpu_launches": 0,
    }


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--peer", type=int, required=True)
    ap.add_argument("--world", type=int, required=True)
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--port", type=int, required=True)
    ap.add_argument("--params", type=int, default=150_000_000)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    a = ap.parse_args()
    print(json.dumps(peer_main(a.peer, a.world, a.device, a.port, a.params, a.steps, a.warmup)), flush=True)
