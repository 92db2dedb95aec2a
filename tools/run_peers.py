#!/usr/bin/env python3
"""Launches a master and N peer_worker processes (one per GPU) and prints their output.
usage: run_peers.py <world> <scenario> [devices comma separated]"""
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "tests"))
import json  # noqa: E402
import subprocess  # noqa: E402

import pccl_b200 as pccl  # noqa: E402

world, scenario = int(sys.argv[1]), sys.argv[2]
devices = sys.argv[3].split(",") if len(sys.argv) > 3 else [f"cuda:{i}" for i in range(world)]
master = pccl.MasterNode("127.0.0.1:0")
master.run()
worker = Path(__file__).resolve().parent.parent / "tests" / "peer_worker.py"
procs = [subprocess.Popen([sys.executable, str(worker), str(master.port), str(r), str(world), devices[r], scenario],
                          stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for r in range(world)]
rc = 0
for r, p in enumerate(procs):
    try:
        out, _ = p.communicate(timeout=int(sys.argv[4]) if len(sys.argv) > 4 else 300)
    except subprocess.TimeoutExpired:
        p.kill()
        out, _ = p.communicate()
        out += "\n[TIMEOUT]"
    print(f"===== rank {r} rc={p.returncode}")
    print(out[-6000:])
    rc |= p.returncode or 0
master.interrupt()
sys.exit(rc)
