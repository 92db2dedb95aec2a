#!/usr/bin/env python3
"""Stand-alone throughput of the three device content hashes (simplehash, CRC32C, mset64) on one GPU.

Every call is `pcclKernelHashSync` (launch, wait, 8-byte read-back), timed per call with CUDA events on the launching
stream after warm-up; the minimum and the median of 30 calls are printed.  1 GiB is far larger than L2 (126 MB), so that
row is an HBM number; the 64 MiB row fits into L2 after the first pass and says so.

  python tools/prof_hashes.py > profiles/round2/hashes_1gpu.json"""
import json
import statistics
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pccl_b200 import ops  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    rows = []
    for nbytes in (64 << 20, 1 << 30):
        x = torch.randint(0, 255, (nbytes,), dtype=torch.uint8, device=dev)
        for name, fn in (("simplehash", ops.simplehash), ("crc32c", ops.crc32c), ("mset64", ops.mset64)):
            for _ in range(3):
                fn(x)
            times = []
            for _ in range(30):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                fn(x)
                e1.record()
                torch.cuda.synchronize()
                times.append(e0.elapsed_time(e1))
            rows.append({"hash": name, "bytes": nbytes, "min_ms": round(min(times), 4), "median_ms": round(statistics.median(times), 4),
                         "GBps_at_min": round(nbytes / min(times) / 1e6, 1), "in_L2": nbytes < (126 << 20)})
    print(json.dumps({"device": torch.cuda.get_device_name(0), "rows": rows}))


if __name__ == "__main__":
    main()
