#!/usr/bin/env python3
"""A peer of the chaos test: joins, then loops  [update_topology]  sync_shared_state  all_reduce  forever,
printing one line per completed reduction.  Survives other peers dying at any point.

Joint calls have to line up across peers: a peer that has just been admitted by the others' update_topology
continues with the call the others make next (sync_shared_state), so the membership calls are skipped in the
first iteration (same structure as the reference's stress peer and DiLoCo examples).

Mode "cross_step" (third argument) is the reference's cross_step_async_reduces_test: ten 128x128 tensors are reduced with
`all_reduce_multiple_with_retry(max_in_flight=8)` on a thread that keeps running while the main thread starts the next
step; it is joined before the next joint call (topology update, shared-state synchronisation) and before the tensors are
refilled."""
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent.parent))
import pccl_b200 as pccl  # noqa: E402

import threading  # noqa: E402

port, device = int(sys.argv[1]), sys.argv[2]
mode = sys.argv[3] if len(sys.argv) > 3 else "basic"
dev = torch.device(device)
if dev.type == "cuda":
    torch.cuda.set_device(dev)
RETRYABLE = (pccl.Result.RANK_CONNECTION_LOST, pccl.Result.TOO_FEW_PEERS, pccl.Result.UPDATE_TOPOLOGY_FAILED, pccl.Result.INVALID_USAGE,
             pccl.Result.INTERNAL_ERROR)


def session():
    comm = pccl.Communicator(f"127.0.0.1:{port}")
    comm.connect(n_attempts=60)
    weights = torch.zeros(4096, device=dev)
    step_t = torch.zeros(1, dtype=torch.int64)
    state = pccl.SharedState([pccl.TensorInfo.from_torch(weights, "w"), pccl.TensorInfo.from_torch(step_t, "step")])
    print("JOINED", flush=True)
    it, errors = 0, 0
    grads = [torch.ones(128, 128, device=dev) for _ in range(10)]
    reducer = None
    outcome = {}

    def reduce_all():
        try:
            descs = [pccl.ReduceOpDescriptor.from_torch(g, g, pccl.ReduceDescriptor(
                count=g.numel(), op=pccl.ReduceOp.SUM, tag=100 + i, operand_descriptor=pccl.ReduceOperandDescriptor(pccl.DataType.FLOAT)))
                     for i, g in enumerate(grads)]
            outcome["info"] = comm.all_reduce_multiple_with_retry(descs, max_in_flight=8)
        except BaseException as e:  # noqa: BLE001 -- handed to the main thread, whatever it is
            outcome["error"] = e

    def join_reducer():
        nonlocal reducer
        if reducer is None:
            return
        reducer.join()
        reducer = None
        if "error" in outcome:
            raise outcome.pop("error")
        info = outcome.pop("info")
        for g in grads:  # whatever the world was when the attempt that succeeded ran, all ten agree with it
            assert float(g[0, 0]) == info.world_size and float(g[-1, -1]) == info.world_size, (float(g[0, 0]), info.world_size)
        print(f"REDUCE {int(step_t)} world={info.world_size} cross-step", flush=True)

    while True:
        try:
            if it > 0:
                if comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < 2 or comm.are_peers_pending():
                    join_reducer()
                    comm.update_topology()
            it += 1
            if comm.get_attribute(pccl.Attribute.PEER_GROUP_WORLD_SIZE) < 2:
                join_reducer()
                time.sleep(0.02)
                continue
            if mode == "cross_step":
                join_reducer()
                comm.sync_shared_state(state)
                for g in grads:
                    g.fill_(1.0)
                weights += 1
                step_t += 1
                state.push_revision()
                errors = 0
                reducer = threading.Thread(target=reduce_all, name="reduce")
                reducer.start()
                time.sleep(0.002)  # "compute" of the next step while the reduction runs
                continue
            comm.sync_shared_state(state)
            x = torch.ones(200_000, device=dev)
            info = comm.all_reduce(x, x, op=pccl.ReduceOp.SUM, tag=0)
            assert float(x[0]) == info.world_size, (float(x[0]), info.world_size)
            weights += 1  # every peer applies the same update, so shared state stays in sync
            step_t += 1
            state.push_revision()
            errors = 0
            print(f"REDUCE {int(step_t)} world={info.world_size}", flush=True)
        except pccl.PCCLError as e:
            if e.result not in RETRYABLE:
                print(f"FATAL {e}", flush=True)
                raise
            errors += 1
            if errors > 50:  # kicked or master link lost: start over with a fresh communicator
                if reducer is not None:
                    reducer.join()
                comm.destroy()
                return
            time.sleep(0.01)


while True:
    session()
