#!/usr/bin/env python3
"""Chaos orchestrator (reference: python/tests/stress_tests/basic_stress_test/stresstest_orchestrator.py): starts a master,
then keeps spawning peers and killing them with SIGKILL at random.  Fails when no reduction completes anywhere
for `--stall` seconds.

  python tests/stress/orchestrator.py --seconds 60 --max-peers 5 [--device cpu]
"""
import argparse
import os
import random
import signal
import subprocess
import sys
import threading
import time
from pathlib import Path

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent.parent))
import pccl_b200 as pccl  # noqa: E402


def run(seconds: float, max_peers: int, stall: float, device: str, seed: int = 0, mode: str = "basic", min_kills: int = 0) -> dict:
    with pccl.MasterNode("127.0.0.1:0") as master:  # leaving the block interrupts the master
        master.run()
        return _run(master, seconds, max_peers, stall, device, seed, mode, min_kills)


def _run(master, seconds: float, max_peers: int, stall: float, device: str, seed: int, mode: str = "basic", min_kills: int = 0) -> dict:
    rng = random.Random(seed)
    port = master.port
    peers = []
    last_progress = [time.time()]
    counts = {"reductions": 0, "spawned": 0, "killed": 0}
    lock = threading.Lock()

    joined = set()  # pids of peers that are in the run (a process spends seconds importing torch before that)

    def watch(p):
        for line in p.stdout:
            if line.startswith("REDUCE"):
                with lock:
                    counts["reductions"] += 1
                    last_progress[0] = time.time()
            elif line.startswith("JOINED"):
                with lock:
                    joined.add(p.pid)
                    if len(joined) == 2:
                        last_progress[0] = time.time()  # only now can anything be reduced: the stall clock starts here
            elif line.startswith("FATAL"):
                counts.setdefault("fatal", []).append(line.strip())

    # device "cuda" = one peer per GPU (kernels of two peers that wait for each other must not share a GPU)
    n_gpus = 0
    if device == "cuda":
        import torch
        n_gpus = torch.cuda.device_count()
        max_peers = min(max_peers, n_gpus)
    gpu_of = {}

    def spawn():
        dev = device
        if device == "cuda":
            used = {gpu_of[q.pid] for q in peers if q.poll() is None}
            free = [g for g in range(n_gpus) if g not in used]
            if not free:
                return
            dev = f"cuda:{free[0]}"
        p = subprocess.Popen([sys.executable, str(HERE / "chaos_peer.py"), str(port), dev, mode], stdout=subprocess.PIPE,
                             stderr=subprocess.DEVNULL, text=True)
        if device == "cuda":
            gpu_of[p.pid] = int(dev.split(":")[1])
        threading.Thread(target=watch, args=(p,), daemon=True).start()
        peers.append(p)
        counts["spawned"] += 1

    spawn()
    spawn()
    # the run proper starts when two peers are in (at most 90 s later: then the stall check below reports it)
    t_wait = time.time() + 90
    while time.time() < t_wait:
        with lock:
            if len(joined) >= 2:
                break
        time.sleep(0.05)
    t_end = time.time() + seconds
    stalled = False
    # (a short run on a slow machine is extended, by a minute at most, until it has seen the kills it was asked to see)
    while time.time() < t_end or (counts["killed"] < min_kills and time.time() < t_end + 60):
        time.sleep(rng.uniform(0.3, 1.2))
        peers[:] = [p for p in peers if p.poll() is None]
        if rng.random() < 0.6:
            if len(peers) < max_peers:
                spawn()
        else:
            # victims are peers that made it into the run, and two of those are left alone: on a loaded machine a process
            # needs longer to start than the pause between two kills, and killing what has not joined yet proves nothing
            with lock:
                in_run = [p for p in peers if p.pid in joined]
            if len(in_run) > 2:
                victim = rng.choice(in_run)
                os.kill(victim.pid, signal.SIGKILL)
                peers.remove(victim)
                counts["killed"] += 1
        with lock:
            started = len(joined) >= 2
        if (started and time.time() - last_progress[0] > stall) or (not started and time.time() > t_wait):
            stalled = True
            break
    for p in peers:
        p.kill()
    master.interrupt()
    counts["stalled"] = stalled
    return counts


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--seconds", type=float, default=60)
    ap.add_argument("--max-peers", type=int, default=5)
    ap.add_argument("--stall", type=float, default=20)
    ap.add_argument("--device", default="cpu")
    ap.add_argument("--mode", default="basic", choices=["basic", "cross_step"])
    a = ap.parse_args()
    out = run(a.seconds, a.max_peers, a.stall, a.device, mode=a.mode)
    print(out)
    sys.exit(1 if out["stalled"] or out.get("fatal") else 0)
