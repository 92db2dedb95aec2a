#!/usr/bin/env python3
"""Variants of the NVLS all-reduce kernel on symmetric buffers under torchrun (>= 4 GPUs): packets in flight per thread,
static against dynamic work distribution, block-wide system fences on/off, SM budget.  One JSON line per (size, config)
from rank 0: kernel time (CUDA events, max over ranks) and block 0's phase times."""
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
if rank == 0:
    master = pccl.MasterNode("127.0.0.1:0")
    master.run()
    box[0] = master.port
dist.broadcast_object_list(box, src=0)
comm = pccl.Communicator(f"127.0.0.1:{box[0]}")
comm.connect(n_attempts=15)
while comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < world:
    comm.update_topology()
sizes = [int(a) for a in sys.argv[1:]] or [1 << 20, 1 << 24, 1 << 26, 600_000_000]
sym = comm.ext_symmetric_empty(max(sizes) // 4, torch.float32)
sym.normal_()
# (label, PCCL_DEVICE_VARIANT, PCCL_DEVICE_MAX_BLOCKS)
configs = [("static_u4", 0, 0), ("dyn_u4", 0x200, 0), ("dyn_u8", 0x202, 0), ("dyn_u16", 0x205, 0), ("dyn_u8_b72", 0x202, 72),
           ("dyn_u8_b148", 0x202, 148), ("static_u4_b148", 0, 148)]
for nbytes in sizes:
    x = sym[:nbytes // 4]
    for label, variant, blocks in configs:
        os.environ["PCCL_DEVICE_VARIANT"] = str(variant)
        os.environ["PCCL_DEVICE_MAX_BLOCKS"] = str(blocks)
        ms, st = [], None
        for it in range(10):
            comm.all_reduce(x, x, op=pccl.ReduceOp.SUM, tag=0)
            st = comm.ext_last_reduce_stats(0)
            if it >= 3:
                ms.append(st["device_ms"])
        ms.sort()
        t = torch.tensor([ms[len(ms) // 2]], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(json.dumps({"bytes": nbytes, "config": label, "plane": st["plane"], "kernel_us": round(float(t.item()) * 1e3, 1),
                              "bus_GBps": round(nbytes * 2 * (world - 1) / world / float(t.item()) / 1e6, 1), "phase_us": st["phase_us"]}), flush=True)
        x.normal_()
comm.destroy()
dist.barrier()
dist.destroy_process_group()
