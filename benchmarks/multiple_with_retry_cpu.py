#!/usr/bin/env python3
"""The reference's DiLoCo call pattern on host tensors: one reduction per parameter tensor of GPT-2 124 M (148 tensors, 498 MB)
through all_reduce_multiple_with_retry, two peers on loopback.  usage: multiple_with_retry_cpu.py ours|reference"""
import sys, time, json, os, subprocess
impl = sys.argv[1]
REPO = str(__import__('pathlib').Path(__file__).resolve().parent.parent)
if len(sys.argv) > 2 and sys.argv[2] == 'peer':
    rank, port = int(sys.argv[3]), int(sys.argv[4])
    if impl == 'reference':
        sys.path.insert(0, REPO + '/baseline/_ref'); import pccl
    else:
        sys.path.insert(0, REPO); import pccl_b200 as pccl
    import torch
    kw = {}
    if impl == 'reference':
        base = 47600 + 16 * rank
        kw = dict(p2p_listen_port=base, shared_state_listen_port=base + 1, benchmark_listen_port=base + 2)
    comm = pccl.Communicator(f"127.0.0.1:{port}", 0, p2p_connection_pool_size=4, **kw)
    comm.connect(n_attempts=30)
    while comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < 2:
        comm.update_topology()
    shapes = [(50257, 768), (1024, 768)]
    for _ in range(12):
        shapes += [(768,), (768,), (768, 2304), (2304,), (768, 768), (768,), (768,), (768,), (768, 3072), (3072,), (3072, 768), (768,)]
    shapes += [(768,), (768,)]
    tensors = [torch.full(s, float(rank + 1)) for s in shapes]
    def run():
        descs = []
        for i, t in enumerate(tensors):
            descs.append(pccl.ReduceOpDescriptor.from_torch(t, t, pccl.ReduceDescriptor(count=t.numel(), op=pccl.ReduceOp.AVG, tag=i,
                         operand_descriptor=pccl.ReduceOperandDescriptor(pccl.DataType.FLOAT), quantization_options=pccl.QuantizationOptions(pccl.DataType.FLOAT, pccl.QuantizationAlgorithm.NONE))))
        t0 = time.perf_counter()
        comm.all_reduce_multiple_with_retry(descs, max_in_flight=16)
        return time.perf_counter() - t0
    run()
    times = [run() for _ in range(3)]
    print("RESULT", json.dumps({"impl": impl, "tensors": len(tensors), "bytes": sum(t.numel() for t in tensors) * 4, "best_s": min(times), "all": times}), flush=True)
    sys.exit(0)
if impl == 'reference':
    sys.path.insert(0, REPO + '/baseline/_ref'); import pccl
else:
    sys.path.insert(0, REPO); import pccl_b200 as pccl
port = 48801 if impl == 'ours' else 48802
m = pccl.MasterNode(f"127.0.0.1:{port}"); m.run()
ps = [subprocess.Popen([sys.executable, __file__, impl, 'peer', str(r), str(port)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True, env=dict(os.environ, OMP_NUM_THREADS="2")) for r in range(2)]
for p in ps:
    out = p.communicate(timeout=900)[0]
    print([l for l in out.splitlines() if l.startswith("RESULT")])
m.interrupt()
