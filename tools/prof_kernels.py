#!/usr/bin/env python3
"""Small driver for profiling single kernels on one GPU (used under ncu and for CUDA-event timing).

  python tools/prof_kernels.py hash      # simplehash + crc32c over 1 GiB
  python tools/prof_kernels.py allreduce # one-rank fused all-reduce over 600 MB
  python tools/prof_kernels.py quant     # block quantisers
  python tools/prof_kernels.py outer     # DiLoCo outer step over 150 M parameters: fused kernel vs three kernels
"""
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import pccl_b200 as pccl  # noqa: E402
from pccl_b200 import ops  # noqa: E402


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def main():
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    dev = torch.device("cuda:0")
    out = {}
    if what in ("hash", "all"):
        for mib in (147, 1024):
            x = torch.randint(0, 2**31 - 1, (mib << 18,), dtype=torch.int32, device=dev)
            import ctypes as C
            from pccl_b200._lib import lib
            scratch = torch.zeros(lib().pcclKernelHashScratchBytes(), dtype=torch.uint8, device=dev)
            res = torch.zeros(1, dtype=torch.int64, device=dev)
            s = torch.cuda.current_stream().cuda_stream
            nbytes = x.numel() * 4
            ms = timeit(lambda: lib().pcclKernelSimpleHash(x.data_ptr(), nbytes, scratch.data_ptr(), res.data_ptr(), s))
            out[f"simplehash_{mib}MiB_GBps"] = round(nbytes / ms / 1e6, 1)
            ms = timeit(lambda: lib().pcclKernelCrc32c(x.data_ptr(), nbytes, scratch.data_ptr(), res.data_ptr(), s))
            out[f"crc32c_{mib}MiB_GBps"] = round(nbytes / ms / 1e6, 1)
            del x
    if what in ("allreduce", "all"):
        n = 150_000_000
        x = torch.randn(n, device=dev)
        y = torch.empty_like(x)
        arena = ops.LocalArena(dev, n * 4)
        ms = timeit(lambda: ops.local_all_reduce(x, y, arena, op=pccl.ReduceOp.AVG))
        out["allreduce_world1_600MB_ms"] = round(ms, 4)
        out["allreduce_world1_600MB_algGBps"] = round(n * 4 / ms / 1e6, 1)
        # HBM traffic of the one-rank path: read src, write stage, read stage, write stage, write dst
        out["allreduce_world1_hbm_GBps"] = round(5 * n * 4 / ms / 1e6, 1)
        for q in (pccl.QuantizationAlgorithm.BLOCK_MIN_MAX, pccl.QuantizationAlgorithm.BLOCK_FP8):
            ms = timeit(lambda: ops.local_all_reduce(x, y, arena, op=pccl.ReduceOp.AVG, quant=q))
            out[f"allreduce_world1_{q.name}_ms"] = round(ms, 4)
    if what in ("quant", "all"):
        n = 150_000_000
        x = torch.randn(n, device=dev)
        for q in (pccl.QuantizationAlgorithm.BLOCK_MIN_MAX, pccl.QuantizationAlgorithm.BLOCK_FP8):
            ms = timeit(lambda: ops.block_quantize(x, q))
            out[f"quantize_{q.name}_GBps_in"] = round(n * 4 / ms / 1e6, 1)
    if what in ("outer", "all"):
        from pccl_b200.parallel.diloco import LocalOuterStep
        n = 150_000_000
        outer = torch.randn(n, device=dev)
        inner = outer - 0.01 * torch.randn(n, device=dev)
        step = LocalOuterStep(outer)
        step.step(inner, want_hash=False)
        out["outer_step_fused_world1_ms"] = round(timeit(lambda: step.step(inner, want_hash=False)), 4)
        # HBM traffic of the one-rank fused kernel: 2N+N staging, N+N reduce, 3N+3N update = 11N
        out["outer_step_fused_world1_hbm_GBps"] = round(11 * n * 4 / out["outer_step_fused_world1_ms"] / 1e6, 1)
        out["outer_step_three_kernels_world1_ms"] = round(timeit(lambda: step.step_three_pass(inner, want_hash=False)), 4)
    print(json.dumps(out))


if __name__ == "__main__":
    main()
