#!/usr/bin/env python3
"""Sweeps kernel tuning knobs of the device all-reduce under torchrun (one communicator, many configs).
Prints one JSON line per config from rank 0."""
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
os.environ["PCCL_DEVICE_TUNE"] = "1"
import pccl_b200 as pccl  # noqa: E402

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
n = int(sys.argv[1]) if len(sys.argv) > 1 else 150_000_000
sym = comm.ext_symmetric_empty(n, torch.float32)
sym.normal_()
reg = torch.randn(n, device=dev)
out = torch.empty_like(reg)
configs = [(1, 144, 64)] + ([(1, b, seg) for b in (72, 144) for seg in (32, 64, 128, 256)] if os.environ.get("PCCL_TUNE_FULL") else [])
for buffers in ("symmetric", "regular"):
    for variant, blocks, seg in (configs[:1] if buffers == "symmetric" else configs):
        os.environ["PCCL_DEVICE_VARIANT"] = str(variant)
        os.environ["PCCL_DEVICE_MAX_BLOCKS"] = str(blocks)
        os.environ["PCCL_PIPELINE_SEGMENT_MB"] = str(seg)
        ms = []
        for it in range(8):
            if buffers == "symmetric":
                comm.all_reduce(sym, sym, op=pccl.ReduceOp.AVG, tag=0)
            else:
                comm.all_reduce(reg, out, op=pccl.ReduceOp.AVG, tag=0)
            st = comm.ext_last_reduce_stats(0)
            if it >= 3:
                ms.append(st["device_ms"])
        t = torch.tensor([sum(ms) / len(ms)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(json.dumps({"buffers": buffers, "variant": variant, "blocks": blocks, "segment_mb": seg, "plane": st["plane"], "kernel_ms": round(float(t.item()), 4),
                              "bus_GBps": round(n * 4 * 2 * (world - 1) / world / float(t.item()) / 1e6, 1)}), flush=True)
    if buffers == "symmetric" and world < 4:
        pass

# warp-specialised in-kernel pipeline: segments sweep, plain all-reduce and the fused outer step
outer_p, inner_p, mom_p = torch.randn(n, device=dev), torch.randn(n, device=dev), torch.zeros(n, device=dev)
for pipe, segs in [(0, 0)] + [(1, k) for k in (3, 4, 5, 6, 8, 12)]:
    os.environ["PCCL_DEVICE_VARIANT"] = "0"
    os.environ["PCCL_DEVICE_MAX_BLOCKS"] = "0"
    os.environ["PCCL_KERNEL_PIPELINE"] = str(pipe)
    os.environ["PCCL_PIPELINE_SEGMENTS"] = str(segs)
    row = {"kernel_pipeline": pipe, "segments": segs}
    for what in ("allreduce", "outer_step"):
        ms = []
        for it in range(8):
            if what == "allreduce":
                comm.all_reduce(reg, out, op=pccl.ReduceOp.AVG, tag=0)
            else:
                comm.ext_diloco_outer_step(outer_p, inner_p, mom_p, lr=0.7, mu=0.9, first=False, tag=0)
            st = comm.ext_last_reduce_stats(0)
            if it >= 3:
                ms.append(st["device_ms"])
        t = torch.tensor([sum(ms) / len(ms)], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        row[what + "_kernel_ms"] = round(float(t.item()), 4)
    if rank == 0:
        print(json.dumps(row), flush=True)
comm.destroy()
dist.barrier()
dist.destroy_process_group()
