#!/usr/bin/env python3
"""Two peers that share ONE GPU (so the NVLink plane is off) all-reduce device tensors over the TCP plane:
the configuration `bench.py --impl reference --gpus 1` runs for the reference.  Prints ms per 600 MB step."""
import json
import subprocess
import sys
import time
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))

if len(sys.argv) > 1 and sys.argv[1] == "peer":
    import torch
    import pccl_b200 as pccl
    port, rank, pool, n = int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5])
    torch.cuda.set_device(0)
    comm = pccl.Communicator(f"127.0.0.1:{port}", p2p_connection_pool_size=pool)
    comm.connect(n_attempts=30)
    while comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < 2:
        comm.update_topology()
        time.sleep(0.005)
    x = torch.randn(n, device="cuda:0")
    for _ in range(3):
        comm.all_reduce(x, x, op=pccl.ReduceOp.AVG)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    k = 5
    for _ in range(k):
        comm.all_reduce(x, x, op=pccl.ReduceOp.AVG)
    e1.record()
    torch.cuda.synchronize()
    print(json.dumps({"rank": rank, "ms": e0.elapsed_time(e1) / k, "plane": comm.ext_last_reduce_stats(0)["plane"]}), flush=True)
    comm.destroy()
else:
    import pccl_b200 as pccl
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 150_000_000
    for pool in (1, 4, 8):
        master = pccl.MasterNode("127.0.0.1:0")
        master.run()
        procs = [subprocess.Popen([sys.executable, __file__, "peer", str(master.port), str(r), str(pool), str(n)], stdout=subprocess.PIPE, text=True)
                 for r in range(2)]
        outs = [json.loads(p.communicate(timeout=600)[0].strip().splitlines()[-1]) for p in procs]
        master.interrupt()
        print(json.dumps({"pool": pool, "ms_per_step": round(max(o["ms"] for o in outs), 2), "plane": outs[0]["plane"],
                          "bus_GBps": round(n * 4 / max(o["ms"] for o in outs) / 1e6, 3)}), flush=True)
