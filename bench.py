#!/usr/bin/env python3
"""Headline benchmark: fp32 AVG all-reduce of DiLoCo pseudo-gradients (150 M parameters by default)
across N peers, one peer per GPU (BASELINE.json: "DiLoCo outer-step fp32 all-reduce of 150M-param
pseudo-gradients", metric "all-reduce bus GB/s ...; DiLoCo outer-step sec").

  python bench.py --gpus 1 --steps K --warmup W
  python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
         bench.py --gpus N --steps K --warmup W
  ... --impl reference      runs the unmodified reference from baseline/_ref (host all-reduce, staged)

What is timed (device time, CUDA events on the launching stream, max over ranks):
  value      K back-to-back all-reduces through the public API (`Communicator.all_reduce` on device
             tensors).  bus GB/s = bytes * 2(P-1)/P / t; with ONE peer the factor is undefined and
             the algorithm bandwidth bytes/t of the one-rank fused kernel is reported instead.
  e2e        the same, but every step first copies that step's pseudo-gradients host->device from
             pinned memory and reads the first reduced element back device->host.
  diloco_outer_step_ms         pseudo-gradient kernel + all-reduce + Nesterov kernel (three launches, symmetric scratch)
  diloco_outer_step_fused_ms   the same step as ONE collective on ordinary tensors (`Communicator.ext_diloco_outer_step`)
  regular_buffers              the headline all-reduce on ordinary torch tensors (staged through the arena)
  nccl_same_run                NCCL all-reduce and the torch + NCCL formulation of the outer step, same process, same timing
  verified_against_nccl        all of the above re-run once on fresh random data and compared with NCCL + torch (untimed)
Inputs (600 MB) are far larger than the 126 MB L2, so no explicit L2 flush is needed between steps.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent
sys.path.insert(0, str(REPO))

PUBLISHED_BASELINE_GBPS = 45.0 / 8.0  # reference docs: 45 Gbit/s sustained all-reduce (docs/md/01_Introduction.md:8)


class ClockSampler:
    """Samples SM clocks and throttle reasons with nvidia-smi while the timed region runs."""

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.samples = []
        self.proc = None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "50", "-i", str(self.gpu)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) >= 6:
                self.samples.append(parts)

    def stop(self) -> dict:
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        for p in self.samples:
            try:
                sm.append(float(p[0]))
                mx.append(float(p[1]))
            except ValueError:
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], p[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": sorted(reasons),
                "samples": len(sm)}


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--params", type=int, default=150_000_000, help="fp32 elements per peer")
    ap.add_argument("--quant", default="none", choices=["none", "minmax_u8", "fp8"])
    ap.add_argument("--sweep", action="store_true", help="also print a size sweep table to stderr")
    ap.add_argument("--buffers", default="symmetric", choices=["symmetric", "regular"],
                    help="symmetric: tensors from Communicator.ext_symmetric_empty (zero-copy); regular: plain torch tensors (staged)")
    return ap.parse_args()


def bus_factor(peers: int) -> float:
    return 2.0 * (peers - 1) / peers if peers > 1 else 1.0


_MASTERS = []


def run_ours(args) -> dict:
    import torch

    import pccl_b200 as pccl
    from pccl_b200 import ops
    from pccl_b200.parallel.diloco import LocalOuterStep, outer_nesterov_, pseudo_gradient

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    # host threads and pinned buffers next to the GPU (matters for the end-to-end number with several peers per box)
    from pccl_b200.utils import bind_to_device
    numa = bind_to_device(local_rank) if os.environ.get("PCCL_BENCH_NUMA", "1") != "0" else {"bound": False, "reason": "disabled"}
    n = args.params
    nbytes = n * 4
    torch.manual_seed(1234 + rank)

    dist = None
    if world > 1:
        import torch.distributed as dist_mod
        dist = dist_mod
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if dist:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if not dist:
            return x
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    grad = torch.randn(n, device=dev)
    out = torch.empty_like(grad)
    pinned = torch.randn(n).pin_memory()
    quant = {"none": None,
             "minmax_u8": pccl.QuantizationOptions(pccl.DataType.UINT8, pccl.QuantizationAlgorithm.MIN_MAX),
             "fp8": pccl.QuantizationOptions(pccl.DataType.FLOAT8_E4M3, pccl.QuantizationAlgorithm.BLOCK_FP8)}[args.quant]

    master = comm = None
    sym = None
    planes = set()
    launches = 0
    dev_ms = [0.0]
    if world > 1:
        port_box = [0]
        if rank == 0:
            master = pccl.MasterNode("127.0.0.1:0")
            _MASTERS.append(master)  # a running master that loses its last reference waits to be interrupted (reference semantics)
            master.run()
            port_box[0] = master.port
        dist.broadcast_object_list(port_box, src=0)
        comm = pccl.Communicator(f"127.0.0.1:{port_box[0]}")
        comm.connect(n_attempts=15)
        while comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < world:
            comm.update_topology()

        enqueue = os.environ.get("PCCL_BENCH_ENQUEUE", "1") != "0"

        def collect_stats():
            nonlocal launches
            comm.ext_wait_stream_ops()
            st = comm.ext_last_reduce_stats(0)
            planes.add(st["plane"] + ("+lease" if st.get("leased") else ""))
            return st

        def reduce_step(src, dst):
            # Stream-ordered, like NCCL's all-reduce: the call returns after the launch (when no vote at the master is
            # needed) and the next one queues up behind it; verdicts are taken in batches and at the end of the region.
            nonlocal launches
            comm.all_reduce(src, dst, op=pccl.ReduceOp.AVG, tag=0, quantization_options=quant, enqueue=enqueue)
            launches += 1
            if not enqueue:
                st = collect_stats()
                dev_ms[0] += st["device_ms"]

        if args.buffers == "symmetric" and quant is None:
            # Peer-visible ("symmetric") buffers from the communicator's arena: the reduction runs in
            # place on them with no staging copy.  This is the registered-buffer path; the plain
            # torch-tensor path is measured as well and reported as `regular_buffers`.
            sym = comm.ext_symmetric_empty(n, torch.float32)
            sym.copy_(grad)
    else:
        arena = ops.LocalArena(dev, nbytes)
        qalgo = {"none": pccl.QuantizationAlgorithm.NONE, "minmax_u8": pccl.QuantizationAlgorithm.BLOCK_MIN_MAX,
                 "fp8": pccl.QuantizationAlgorithm.BLOCK_FP8}[args.quant]

        def reduce_step(src, dst):
            nonlocal launches
            ops.local_all_reduce(src, dst, arena, op=pccl.ReduceOp.AVG, quant=qalgo)
            planes.add("p2p_two_shot(world=1)")
            launches += 1

    last_timing = {}

    def timed(fn, steps):
        """Every step between its own pair of CUDA events on the launching stream; the whole region bracketed by a
        barrier + synchronize.  Returns the median step (max over ranks); mean, p95 and max are kept in last_timing
        so that one stalled step shows up instead of silently shifting the number."""
        barrier()
        evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
        evs[0].record()
        for i in range(steps):
            fn()
            evs[i + 1].record()
        if comm is not None:
            comm.ext_wait_stream_ops()  # verdicts of everything enqueued in the region
        torch.cuda.synchronize()
        per = sorted(evs[i].elapsed_time(evs[i + 1]) for i in range(steps))
        total = evs[0].elapsed_time(evs[steps])
        barrier()
        med = per[len(per) // 2] if len(per) % 2 else 0.5 * (per[len(per) // 2 - 1] + per[len(per) // 2])
        p95 = per[min(len(per) - 1, int(0.95 * len(per)))]
        last_timing.clear()
        last_timing.update({"median": max_over_ranks(med), "mean": max_over_ranks(total / steps), "p95": max_over_ranks(p95),
                            "max": max_over_ranks(per[-1]), "min": max_over_ranks(per[0])})
        return last_timing["median"]

    # ---- kernel-level number -----------------------------------------------------------------
    main_src, main_dst = (sym, sym) if sym is not None else (grad, out)
    for _ in range(max(3, args.warmup)):
        reduce_step(main_src, main_dst)
    launches = 0
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    dev_ms[0] = 0.0
    ms_step = timed(lambda: reduce_step(main_src, main_dst), args.steps)
    main_timing = {k: round(v, 4) for k, v in last_timing.items()}
    if comm is not None:
        st_last = collect_stats()
        if not dev_ms[0]:
            dev_ms[0] = st_last["device_ms"] * args.steps  # enqueued operations: the last one's kernel time stands for all
    main_planes = sorted(planes)
    kernel_ms = max_over_ranks(dev_ms[0] / args.steps) if dev_ms[0] else None
    timed_launches = launches

    # ---- end to end: H2D of the step's input, all-reduce, D2H of a result scalar ------------------
    def e2e_step():
        main_src.copy_(pinned, non_blocking=True)
        reduce_step(main_src, main_dst)
        return float(main_dst[0].item())  # device-to-host read of the result: waits for the stream, i.e. for the reduction

    for _ in range(3):
        e2e_step()
    ms_e2e = timed(e2e_step, max(3, args.steps // 2))
    clocks = sampler.stop() if rank == 0 else {}  # sampled across both timed regions

    # ---- DiLoCo outer step ------------------------------------------------------------------------
    outer = torch.randn(n, device=dev)
    inner = outer - 0.01 * torch.randn(n, device=dev)
    buf = torch.zeros_like(outer)
    pg = sym if sym is not None else torch.empty_like(outer)
    state = {"k": 0}

    def outer_step():
        pseudo_gradient(pg, outer, inner)
        reduce_step(pg, pg if world > 1 else out)
        outer_nesterov_(outer, buf, pg if world > 1 else out, inner, 0.7, 0.9, state["k"] == 0)
        state["k"] += 1

    for _ in range(3):
        outer_step()
    ms_outer = timed(outer_step, max(3, args.steps // 2))

    # the same step as ONE collective on plain torch tensors (one kernel per peer: pseudo-gradient while
    # staging, reduction over NVLink, Nesterov update while writing out)
    fused_planes = set()
    if world > 1:
        def fused_outer_step():
            comm.ext_diloco_outer_step(outer, inner, buf, lr=0.7, mu=0.9, first=False, tag=0)
            fused_planes.add(comm.ext_last_reduce_stats(0)["plane"])
    else:
        def fused_outer_step():
            ops.local_outer_step(outer, inner, buf, arena, 0.7, 0.9, False)
    for _ in range(3):
        fused_outer_step()
    ms_outer_fused = timed(fused_outer_step, max(3, args.steps // 2))
    ops.simplehash(outer)  # first call creates the per-device scratch and stream
    t0 = time.perf_counter()
    digest = ops.simplehash(outer)
    ms_hash = (time.perf_counter() - t0) * 1e3

    # ---- the same reduction on plain torch tensors (staged through the arena) ----------------------
    regular = None
    if sym is not None:
        planes.clear()
        for _ in range(3):
            reduce_step(grad, out)
        dev_ms[0] = 0.0
        ms_reg = timed(lambda: reduce_step(grad, out), args.steps)
        st_last = collect_stats()
        if not dev_ms[0]:
            dev_ms[0] = st_last["device_ms"] * args.steps
        regular = {"ms_per_step": round(ms_reg, 4), "bus_GBps": round(nbytes * bus_factor(world) / (ms_reg * 1e-3) / 1e9, 2),
                   "kernel_ms_per_step": round(max_over_ranks(dev_ms[0] / args.steps), 4), "planes": sorted(planes)}

    # ---- NCCL on the same tensors, same timing (the baseline the reference's own examples use) ----
    nccl = None
    if dist:
        for _ in range(3):
            dist.all_reduce(grad, op=dist.ReduceOp.AVG)
        ms_nccl = timed(lambda: dist.all_reduce(grad, op=dist.ReduceOp.AVG), args.steps)
        nccl = {"ms_per_step": round(ms_nccl, 4), "bus_GBps": round(nbytes * bus_factor(world) / (ms_nccl * 1e-3) / 1e9, 2)}
        # and the outer step the way a torch + NCCL program writes it (sync_diloco.py:480-530 with NCCL as the transport)
        tmp = torch.empty_like(outer)

        def nccl_outer_step():
            torch.sub(outer, inner, out=grad)
            dist.all_reduce(grad, op=dist.ReduceOp.AVG)
            buf.mul_(0.9).add_(grad)
            torch.add(grad, buf, alpha=0.9, out=tmp)
            outer.add_(tmp, alpha=-0.7)
            inner.copy_(outer)

        for _ in range(3):
            nccl_outer_step()
        nccl["diloco_outer_step_ms"] = round(timed(nccl_outer_step, max(3, args.steps // 2)), 4)
        del tmp

    # ---- self-check (outside every timed region): our results against NCCL + torch on fresh random data -------------
    verified = None
    if dist and quant is None:
        x = torch.randn(n, device=dev)
        want = x.clone()
        dist.all_reduce(want, op=dist.ReduceOp.AVG)
        reduce_step(x, out)
        ok = torch.allclose(out, want, rtol=1e-5, atol=1e-6)
        if sym is not None:
            sym.copy_(x)
            reduce_step(sym, sym)
            ok = ok and torch.allclose(sym, want, rtol=1e-5, atol=1e-6)
        p = torch.randn(n, device=dev)
        dist.broadcast(p, src=0)                       # outer parameters are replicated
        q = p - 0.01 * torch.randn(n, device=dev)      # every peer drifted differently
        m = torch.randn(n, device=dev)
        dist.broadcast(m, src=0)
        g = p - q
        dist.all_reduce(g, op=dist.ReduceOp.AVG)
        m_want = 0.9 * m + g
        p_want = p - 0.7 * (g + 0.9 * m_want)
        comm.ext_diloco_outer_step(p, q, m, lr=0.7, mu=0.9, first=False, tag=0)
        ok = ok and torch.allclose(p, p_want, rtol=1e-4, atol=1e-5) and torch.allclose(m, m_want, rtol=1e-4, atol=1e-5) and torch.equal(p, q)
        flag = torch.tensor([1.0 if ok else 0.0], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        verified = bool(flag.item() == 1.0)
        del x, want, p, q, m, g, m_want, p_want

    factor = bus_factor(world)
    value = nbytes * factor / (ms_step * 1e-3) / 1e9
    e2e_value = nbytes * factor / (ms_e2e * 1e-3) / 1e9
    result = {
        "metric": "allreduce_bus_GBps",
        "value": round(value, 3), "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": max(3, args.warmup),
        "ms_per_step": round(ms_step, 4), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": round(value / PUBLISHED_BASELINE_GBPS, 2), "dtype": "fp32", "data": "synthetic",
        "impl": "ours",
        "config": {"model": "DiLoCo pseudo-gradients, GPT-2-class 150M params fp32", "params": n, "global_batch": None, "seq_len": None,
                   "parallelism": f"dp{world}", "op": "avg", "quantization": args.quant, "planes": main_planes,
                   "buffers": "symmetric (Communicator.ext_symmetric_empty, in place)" if sym is not None else "regular torch tensors",
                   "l2_policy": "inputs (600 MB) larger than L2 (126 MB); no explicit flush",
                   "bus_factor": factor, "numa": numa,
                   "note": "one peer: the public API refuses to reduce alone (pcclTooFewPeers, as the reference), so the same fused kernel runs "
                           "with world size 1 through ops.local_all_reduce; bus factor 1" if world == 1 else
                           "bus GB/s = bytes*2(P-1)/P/t; t = median step (device time, max over ranks) incl. control plane"},
        "e2e": {"value": round(e2e_value, 3), "unit": "GB/s", "ms_per_step": round(ms_e2e, 4), "h2d_bytes_per_step": nbytes, "d2h_bytes_per_step": 4},
        "gpu_launches": timed_launches, "api": "Communicator.all_reduce(..., enqueue=%s), stream-ordered" % (world > 1 and os.environ.get("PCCL_BENCH_ENQUEUE", "1") != "0"),
        "step_ms": main_timing,
        "call_overhead_ms": round(ms_step - kernel_ms, 4) if kernel_ms else None,
        "diloco_outer_step_ms": round(ms_outer, 4),
        "diloco_outer_step_fused_ms": round(ms_outer_fused, 4), "diloco_outer_step_fused_planes": sorted(fused_planes), "shared_state_hash_ms": round(ms_hash, 3), "shared_state_hash": digest,
        "clocks": clocks,
        "kernel_ms_per_step": round(kernel_ms, 4) if kernel_ms else None,
        "regular_buffers": regular,
        "nccl_same_run": nccl,
        "verified_against_nccl": verified,
    }
    if comm is not None:
        comm.destroy()
    if dist:
        barrier()
    if master is not None:
        master.interrupt()
    if dist:
        dist.destroy_process_group()
    return result if rank == 0 else {}


def run_reference(args) -> dict:
    ref_dir = REPO / "baseline" / "_ref"
    marker = ref_dir / "pccl"
    if not marker.exists():
        # normally done by __graft_entry__.build(); try once here when the sources are on this machine
        from baseline.install_reference import REF_SRC, install, installed
        if REF_SRC.exists():
            if int(os.environ.get("RANK", "0")) == 0:
                install(verbose=False)
            else:
                deadline = time.time() + 300
                while not installed() and time.time() < deadline:
                    time.sleep(1)
    if not marker.exists():
        return {"impl": "reference", "unavailable": "reference is not installed in baseline/_ref and could not be built here (baseline/install_reference.py)"}
    from baseline.reference_arm import run as run_ref  # noqa: WPS433
    return run_ref(args)


def main():
    args = parse_args()
    out = run_reference(args) if args.impl == "reference" else run_ours(args)
    if out:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
