#!/usr/bin/env python3
"""Host tensors over the TCP plane on one machine: this library against the unmodified reference build (baseline/_ref).

The reference's only data path is this one (host memory, ring over sockets; ccoip/src/cpp/reduce.cpp:528-784), and it is
the path every multi-host run of this library uses between machines.  Needs no GPU: wall-clock time around blocking
`all_reduce` calls, median of the repetitions, maximum over peers; bus GB/s = bytes * 2(P-1)/P / t.

  python benchmarks/tcp_ring_cpu.py --peers 2 4 --sizes 1024 1048576 67108864 --impl ours reference > profiles/round2/tcp_ring_cpu.jsonl

One process per peer (spawned by this script), one master per (implementation, peer count).  fp32 SUM, plus uint8 min-max
quantisation for the large sizes."""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
REF = REPO / "baseline" / "_ref"
STATE_TENSORS = 1


def load(impl: str):
    if impl == "reference":
        sys.path.insert(0, str(REF))
        import pccl
        return pccl
    sys.path.insert(0, str(REPO))
    import pccl_b200
    return pccl_b200


def peer(args) -> None:
    import numpy as np
    pccl = load(args.impl)
    base = 47000 + 16 * args.rank + (400 if args.impl == "reference" else 0)
    kw = dict(p2p_connection_pool_size=args.pool)
    if args.impl == "reference":
        kw.update(p2p_listen_port=base, shared_state_listen_port=base + 1, benchmark_listen_port=base + 2)
    comm = pccl.Communicator(f"127.0.0.1:{args.port}", 0, **kw)
    comm.connect(n_attempts=30)
    while comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < args.world:
        comm.update_topology()
    rows = []
    for size in args.sizes:
        n = max(1, size // 4)
        x = np.random.default_rng(args.rank).standard_normal(n).astype(np.float32)
        y = np.empty_like(x)
        for quant in (None, "u8") if size >= (1 << 20) else (None,):
            q = None
            if quant:
                q = pccl.QuantizationOptions(pccl.DataType.UINT8, pccl.QuantizationAlgorithm.MIN_MAX)
            reps = max(3, min(200, int(2e8 // max(size, 1 << 16))))
            times = []
            for i in range(reps + 2):
                t0 = time.perf_counter()
                if q is None:
                    comm.all_reduce(x, y, op=pccl.ReduceOp.SUM, tag=0)
                else:
                    comm.all_reduce(x, y, op=pccl.ReduceOp.SUM, tag=0, quantization_options=q)
                if i >= 2:
                    times.append(time.perf_counter() - t0)
            rows.append({"bytes": n * 4, "quant": quant or "none", "median_s": statistics.median(times), "min_s": min(times), "reps": reps})
    # shared state: rank 0 holds the content everybody must end up with (two peers of the majority would be needed for an
    # election; SEND_ONLY / RECEIVE_ONLY states it directly), then a second synchronisation in which nothing differs
    revision = 0  # one counter for the whole run: the master expects every synchronisation to carry the next number
    for size in args.state_sizes:
        n = max(1, size // 4)
        w = np.full(n, 1.0 + (0 if args.rank == 0 else args.rank), dtype=np.float32)
        parts = np.array_split(w, args.state_tensors)  # views: a model is many tensors, not one
        state = pccl.SharedState([pccl.TensorInfo.from_numpy(p, f"w{i}") for i, p in enumerate(parts)])
        state.revision = revision
        strategy = pccl.SharedStateSyncStrategy.SEND_ONLY if args.rank == 0 else pccl.SharedStateSyncStrategy.RECEIVE_ONLY
        t0 = time.perf_counter()
        info = comm.sync_shared_state(state, strategy)
        t_move = time.perf_counter() - t0
        assert float(w[0]) == 1.0 and float(w[-1]) == 1.0
        state.revision = revision + 1
        t0 = time.perf_counter()
        info2 = comm.sync_shared_state(state)
        t_check = time.perf_counter() - t0
        revision += 2
        assert info2.tx_bytes == 0 and info2.rx_bytes == 0
        rows.append({"bytes": n * 4, "quant": "shared_state", "median_s": t_move, "min_s": t_check, "reps": 1,
                     "moved": int(info.tx_bytes + info.rx_bytes)})
    print("ROWS " + json.dumps(rows), flush=True)
    if hasattr(comm, "destroy"):
        comm.destroy()


def run(impl: str, world: int, sizes, pool: int, state_sizes=()):
    pccl = load(impl)
    port = 48700 + world + (50 if impl == "reference" else 0)
    master = pccl.MasterNode(f"127.0.0.1:{port}")
    master.run()
    env = dict(os.environ, OMP_NUM_THREADS="2")
    procs = [subprocess.Popen([sys.executable, __file__, "--as-peer", "--impl", impl, "--rank", str(r), "--world", str(world), "--port", str(port),
                               "--pool", str(pool), "--state-tensors", str(STATE_TENSORS), "--state-sizes", *map(str, state_sizes), "--sizes", *map(str, sizes)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, env=env)
             for r in range(world)]
    per_peer = []
    for p in procs:
        out, _ = p.communicate(timeout=3600)
        line = [l for l in out.splitlines() if l.startswith("ROWS ")]
        assert p.returncode == 0 and line, out[-2000:]
        per_peer.append(json.loads(line[-1][5:]))
    master.interrupt()
    for i, row in enumerate(per_peer[0]):
        t = max(pp[i]["median_s"] for pp in per_peer)
        if row["quant"] == "shared_state":
            print(json.dumps({"impl": impl, "peers": world, "bytes": row["bytes"], "what": "shared_state", "sync_ms": round(t * 1e3, 3),
                              "GBps": round(row["bytes"] / t / 1e9, 3), "recheck_ms": round(max(pp[i]["min_s"] for pp in per_peer) * 1e3, 3)}), flush=True)
            continue
        print(json.dumps({"impl": impl, "peers": world, "bytes": row["bytes"], "quant": row["quant"], "ms": round(t * 1e3, 4),
                          "bus_GBps": round(row["bytes"] * 2 * (world - 1) / world / t / 1e9, 3), "reps": row["reps"], "pool": pool}), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--impl", nargs="+", default=["ours", "reference"])
    ap.add_argument("--peers", nargs="+", type=int, default=[2, 4])
    ap.add_argument("--sizes", nargs="+", type=int, default=[1 << 10, 1 << 16, 1 << 20, 1 << 24, 1 << 26, 1 << 28])
    ap.add_argument("--state-sizes", nargs="*", type=int, default=[])
    ap.add_argument("--state-tensors", type=int, default=1)
    ap.add_argument("--pool", type=int, default=4)
    ap.add_argument("--as-peer", action="store_true")
    ap.add_argument("--rank", type=int, default=0)
    ap.add_argument("--world", type=int, default=2)
    ap.add_argument("--port", type=int, default=0)
    args = ap.parse_args()
    global STATE_TENSORS
    STATE_TENSORS = args.state_tensors
    if args.as_peer:
        args.impl = args.impl[0]
        peer(args)
        return
    for impl in args.impl:
        if impl == "reference" and not (REF / "pccl").exists():
            print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref is not installed (python baseline/install_reference.py)"}))
            continue
        for world in args.peers:
            run(impl, world, args.sizes, args.pool, args.state_sizes)


if __name__ == "__main__":
    main()
