#!/usr/bin/env python3
"""NVLink byte counters around a loop of 600 MB all-reduces (under torchrun): how many bytes each GPU really sends and
receives per operation, against the algorithm's expectation (NVLS: N(1+1/P) per direction, two-shot: 2N(P-1)/P).
Counters come from `nvidia-smi nvlink -gt d` (per link, KiB); every rank reads its own GPU.  One JSON line from rank 0."""
import json
import os
import re
import subprocess
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import pccl_b200 as pccl  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
box = [0]
if rank == 0:
    master = pccl.MasterNode("127.0.0.1:0")
    master.run()
    box[0] = master.port
dist.broadcast_object_list(box, src=0)
comm = pccl.Communicator(f"127.0.0.1:{box[0]}")
comm.connect(n_attempts=15)
while comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < world:
    comm.update_topology()


def counters():
    try:
        out = subprocess.run(["nvidia-smi", "nvlink", "-gt", "d", "-i", str(local)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True,
                             timeout=20).stdout
    except Exception:
        return None
    tx = sum(int(m) for m in re.findall(r"Data Tx:\s*(\d+)\s*KiB", out))
    rx = sum(int(m) for m in re.findall(r"Data Rx:\s*(\d+)\s*KiB", out))
    return (tx * 1024, rx * 1024) if (tx or rx) else None


n = int(sys.argv[1]) if len(sys.argv) > 1 else 150_000_000
iters = 100
sym = comm.ext_symmetric_empty(n, torch.float32)
sym.normal_()
reg = torch.randn(n, device=dev)
rows = []
for label, buf in (("symmetric", sym), ("regular", reg), ("nccl", reg)):
    def step():
        if label == "nccl":
            dist.all_reduce(buf, op=dist.ReduceOp.SUM)
        else:
            comm.all_reduce(buf, buf, op=pccl.ReduceOp.SUM, tag=0, enqueue=True)
    for _ in range(5):
        step()
    if label != "nccl":
        comm.ext_wait_stream_ops()
    dist.barrier()
    torch.cuda.synchronize()
    before = counters()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        step()
    if label != "nccl":
        comm.ext_wait_stream_ops()
    e1.record()
    torch.cuda.synchronize()
    dist.barrier()
    after = counters()
    ms = e0.elapsed_time(e1) / iters
    if before and after:
        tx, rx = (after[0] - before[0]) / iters, (after[1] - before[1]) / iters
    else:
        tx = rx = float("nan")
    t = torch.tensor([ms, tx, rx], device=dev, dtype=torch.float64)
    gathered = [torch.zeros_like(t) for _ in range(world)]
    dist.all_gather(gathered, t)
    plane = comm.ext_last_reduce_stats(0)["plane"] if label != "nccl" else "nccl"
    rows.append({"impl": label, "plane": plane, "bytes": n * 4, "world": world, "ms": round(max(float(g[0]) for g in gathered), 4),
                 "tx_bytes_per_op_per_gpu": [round(float(g[1])) if g[1] == g[1] else None for g in gathered],
                 "rx_bytes_per_op_per_gpu": [round(float(g[2])) if g[2] == g[2] else None for g in gathered],
                 "expected_nvls_bytes_per_direction": round(n * 4 * (1 + 1 / world)),
                 "expected_two_shot_bytes_per_direction": round(n * 4 * 2 * (world - 1) / world)})
if rank == 0:
    for r in rows:
        print(json.dumps(r), flush=True)
comm.destroy()
dist.barrier()
dist.destroy_process_group()
