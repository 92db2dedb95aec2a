#!/usr/bin/env python3
"""All-reduce sweep 1 KB - 1 GB (BASELINE.json config 5) under torchrun: this library (regular and
symmetric buffers, 1 / 4 / 16 concurrent tags) against NCCL on the same box, device-timed, max over ranks.

  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 benchmarks/sweep.py [--quant]
Prints one JSON line per (size, variant) from rank 0.
"""
import argparse
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import pccl_b200 as pccl  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--max-mb", type=int, default=1024)
ap.add_argument("--iters", type=int, default=10)
ap.add_argument("--quant", action="store_true", help="also run the quantised variants (fp32 only)")
ap.add_argument("--concurrency", action="store_true", help="also run 4 and 16 concurrent tags")
args = ap.parse_args()

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


def timed(fn, iters):
    for _ in range(3):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def emit(**kw):
    if rank == 0:
        print(json.dumps(kw), flush=True)


factor = 2.0 * (world - 1) / world
sizes = [1 << k for k in range(10, 31, 2)] + [600_000_000]
sizes = sorted(s for s in set(sizes) if s <= args.max_mb << 20)
sym_all = comm.ext_symmetric_empty((min(max(sizes), 1 << 30)) // 4, torch.float32)


def ours(src, dst, enqueue, tag=0, q=None):
    def step():
        comm.all_reduce(src, dst, op=pccl.ReduceOp.SUM, tag=tag, quantization_options=q, enqueue=enqueue)
    return step


def timed_ours(src, dst, iters, enqueue, q=None):
    """enqueue=False: every call returns when its reduction is done (the reference's contract; what a caller that needs the
    result on the host sees).  enqueue=True: stream-ordered like NCCL -- calls queue up, one wait at the end (inside the
    timed region)."""
    step = ours(src, dst, enqueue, q=q)

    def run():
        step()
    if not enqueue:
        return timed(run, iters)
    for _ in range(3):
        step()
    comm.ext_wait_stream_ops()
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        step()
    comm.ext_wait_stream_ops()
    e1.record()
    torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / iters], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


for nbytes in sizes:
    n = nbytes // 4
    x = torch.randn(n, device=dev)
    y = torch.empty_like(x)
    iters = args.iters if nbytes >= (1 << 24) else args.iters * 5
    for buffers, (src, dst) in (("regular", (x, y)), ("symmetric", (sym_all[:n], sym_all[:n]))):
        for enqueue in (False, True):
            ms = timed_ours(src, dst, iters, enqueue)
            st = comm.ext_last_reduce_stats(0)
            emit(impl="ours", buffers=buffers, api="enqueue" if enqueue else "blocking", bytes=nbytes, world=world, ms=round(ms, 5),
                 bus_GBps=round(nbytes * factor / ms / 1e6, 2), plane=st["plane"], leased=st["leased"], kernel_us=round(st["device_ms"] * 1e3, 1),
                 phase_us=st["phase_us"], tags=1)
    ms = timed(lambda: dist.all_reduce(x, op=dist.ReduceOp.SUM), iters)
    emit(impl="nccl", buffers="regular", api="enqueue", bytes=nbytes, world=world, ms=round(ms, 5), bus_GBps=round(nbytes * factor / ms / 1e6, 2), plane="nccl", tags=1)
    if args.quant and nbytes >= (1 << 20):
        want = x.clone()
        dist.all_reduce(want, op=dist.ReduceOp.SUM)
        for name, q in (("u8", pccl.QuantizationOptions(pccl.DataType.UINT8, pccl.QuantizationAlgorithm.MIN_MAX)),
                        ("fp8", pccl.QuantizationOptions(pccl.DataType.FLOAT8_E4M3, pccl.QuantizationAlgorithm.BLOCK_FP8))):
            ms = timed_ours(x, y, iters, True, q=q)
            err = (y - want).abs().max().item() / max(want.abs().max().item(), 1e-30)
            rms = ((y - want).pow(2).mean().sqrt() / want.pow(2).mean().sqrt()).item()
            emit(impl="ours", buffers="regular", api="enqueue", quant=name, bytes=nbytes, world=world, ms=round(ms, 5),
                 bus_GBps_fp32_equiv=round(nbytes * factor / ms / 1e6, 2), plane=comm.ext_last_reduce_stats(0)["plane"], tags=1,
                 max_abs_err_over_max=round(err, 5), rel_rms_err=round(rms, 5))
        del want
    if args.concurrency and nbytes <= (1 << 26):
        for tags in (4, 16):
            xs = [torch.randn(n, device=dev) for _ in range(tags)]

            def many_async():
                hs = [comm.all_reduce_async(t, t, op=pccl.ReduceOp.SUM, tag=k) for k, t in enumerate(xs)]
                for h in hs:
                    assert h.wait()[0]

            def many_enqueued():
                for k, t in enumerate(xs):
                    comm.all_reduce(t, t, op=pccl.ReduceOp.SUM, tag=k, enqueue=True)
                comm.ext_wait_stream_ops()
            for api, fn in (("async", many_async), ("enqueue", many_enqueued)):
                ms = timed(fn, max(3, iters // 2))
                emit(impl="ours", buffers="regular", api=api, bytes=nbytes, world=world, ms=round(ms, 5),
                     bus_GBps=round(tags * nbytes * factor / ms / 1e6, 2), plane=comm.ext_last_reduce_stats(0)["plane"], tags=tags)
            del xs
    del x, y
comm.destroy()
dist.barrier()
dist.destroy_process_group()
