#!/usr/bin/env python3
"""One-rank staged all-reduce of 150 M fp32 (the kernel bench.py times at N=1), a few launches, for ncu."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import pccl_b200 as pccl  # noqa: E402
from pccl_b200 import ops  # noqa: E402

n = 150_000_000
dev = torch.device("cuda:0")
x = torch.randn(n, device=dev)
y = torch.empty_like(x)
arena = ops.LocalArena(dev, n * 4)
for _ in range(6):
    ops.local_all_reduce(x, y, arena, op=pccl.ReduceOp.AVG)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    ops.local_all_reduce(x, y, arena, op=pccl.ReduceOp.AVG)
e1.record()
torch.cuda.synchronize()
print("ms per launch", e0.elapsed_time(e1) / 10, "equal", bool(torch.equal(x, y)))
