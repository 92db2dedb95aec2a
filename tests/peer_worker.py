#!/usr/bin/env python3
"""One peer of a multi-process test.  Usage: peer_worker.py <master_port> <rank> <world> <device|cpu> <scenario>

Prints one JSON line with per-case results; exit code 0 when every case matched its expectation."""
import json
import sys
import time
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import pccl_b200 as pccl  # noqa: E402
from pccl_b200 import ops  # noqa: E402


def make_input(case_id: int, rank: int, n: int, dtype: torch.dtype) -> torch.Tensor:
    g = torch.Generator().manual_seed(1000 * case_id + rank)
    if dtype.is_floating_point:
        return (torch.randn(n, generator=g) * 2).to(dtype)
    return torch.randint(-50, 50, (n,), generator=g).to(dtype)


def expected(case_id: int, world: int, n: int, dtype: torch.dtype, op: pccl.ReduceOp) -> torch.Tensor:
    acc_dtype = torch.float64 if dtype == torch.float64 else (torch.float32 if dtype.is_floating_point else torch.int64)
    acc = make_input(case_id, 0, n, dtype).to(acc_dtype)
    for r in range(1, world):
        x = make_input(case_id, r, n, dtype).to(acc_dtype)
        if op in (pccl.ReduceOp.SUM, pccl.ReduceOp.AVG):
            acc = acc + x
        elif op == pccl.ReduceOp.MAX:
            acc = torch.maximum(acc, x)
        elif op == pccl.ReduceOp.MIN:
            acc = torch.minimum(acc, x)
        else:
            acc = acc * x
    if op == pccl.ReduceOp.AVG:
        acc = acc / world if dtype.is_floating_point else torch.div(acc, world, rounding_mode="trunc")
    return acc


CASES = [
    # id, n, dtype, op, quant, in_place, unaligned
    (1, 1000, torch.float32, pccl.ReduceOp.SUM, None, False, False),
    (2, 10_000_019, torch.float32, pccl.ReduceOp.AVG, None, False, False),
    (3, 1_000_003, torch.bfloat16, pccl.ReduceOp.SUM, None, False, False),
    (4, 70_001, torch.int32, pccl.ReduceOp.MAX, None, False, False),
    (5, 500_000, torch.float64, pccl.ReduceOp.MIN, None, False, False),
    (6, 3_000_000, torch.float32, pccl.ReduceOp.SUM, None, True, False),
    (7, 2_000_000, torch.float32, pccl.ReduceOp.SUM, "u8", False, False),
    (8, 2_000_001, torch.float32, pccl.ReduceOp.AVG, "fp8", False, False),
    (9, 400_005, torch.float16, pccl.ReduceOp.SUM, None, False, True),
    (10, 3, torch.float32, pccl.ReduceOp.SUM, None, False, False),
    (11, 90_000_000, torch.float32, pccl.ReduceOp.AVG, None, True, False),  # 360 MB: more than one staging round
    (12, 1_000_000, torch.int64, pccl.ReduceOp.AVG, None, False, False),
]


def main():
    port, rank, world, device, scenario = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4], sys.argv[5]
    dev = torch.device(device)
    if dev.type == "cuda":
        torch.cuda.set_device(dev)
    comm = pccl.Communicator(f"127.0.0.1:{port}")
    comm.connect(n_attempts=20)
    while comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < world:
        comm.update_topology()
        time.sleep(0.005)
    results = {"rank": rank, "cases": []}
    ok_all = True
    if scenario == "allreduce":
        for cid, n, dtype, op, quant, in_place, unaligned in CASES:
            x = make_input(cid, rank, n + (1 if unaligned else 0), dtype)
            if unaligned:
                # a view that starts one element into the allocation (not 16-byte aligned)
                x = make_input(cid, rank, n, dtype)
                store = torch.empty(n + 1, dtype=dtype, device=dev)
                send = store[1:]
                send.copy_(x)
                recv_store = torch.empty(n + 1, dtype=dtype, device=dev)
                recv = recv_store[1:]
            else:
                send = x.to(dev)
                recv = send if in_place else torch.empty_like(send)
            qopt = None
            if quant == "u8":
                qopt = pccl.QuantizationOptions(pccl.DataType.UINT8, pccl.QuantizationAlgorithm.MIN_MAX)
            elif quant == "fp8":
                qopt = pccl.QuantizationOptions(pccl.DataType.FLOAT8_E4M3, pccl.QuantizationAlgorithm.BLOCK_FP8)
            info = comm.all_reduce(send, recv, op=op, tag=cid, quantization_options=qopt)
            st = comm.ext_last_reduce_stats(cid)
            want = expected(cid, world, n, dtype, op)
            got = recv.cpu()
            if quant:
                scale = want.abs().max().item()
                tol = (0.03 if quant == "u8" else 0.12) * scale * (1 + 0.3 * world)
                ok = (got.double() - want.double()).abs().max().item() <= tol
            elif dtype.is_floating_point:
                rtol = {torch.float32: 1e-5, torch.float64: 1e-12, torch.bfloat16: 3e-2, torch.float16: 5e-3}[dtype]
                ok = torch.allclose(got.double(), want.double(), rtol=rtol, atol=rtol * 8)
            else:
                ok = torch.equal(got.to(torch.int64), want)
            digest = ops.simplehash(recv if not unaligned else recv.clone())
            ok = ok and info.world_size == world
            ok_all = ok_all and ok
            results["cases"].append({"id": cid, "ok": bool(ok), "plane": st["plane"], "ms": round(st["device_ms"], 4), "hash": digest,
                                     "launches": st["kernel_launches"]})
    elif scenario == "concurrent":
        n = 4_000_000
        a = make_input(101, rank, n, torch.float32).to(dev)
        b = make_input(102, rank, n, torch.float32).to(dev)
        ra, rb = torch.empty_like(a), torch.empty_like(b)
        h1 = comm.all_reduce_async(a, ra, op=pccl.ReduceOp.SUM, tag=1)
        h2 = comm.all_reduce_async(b, rb, op=pccl.ReduceOp.MAX, tag=2)
        ok1, _, _ = h1.wait()
        ok2, _, _ = h2.wait()
        wa = expected(101, world, n, torch.float32, pccl.ReduceOp.SUM)
        wb = expected(102, world, n, torch.float32, pccl.ReduceOp.MAX)
        ok_all = ok1 and ok2 and torch.allclose(ra.cpu(), wa, rtol=1e-5, atol=1e-5) and torch.equal(rb.cpu(), wb)
        results["cases"].append({"id": "concurrent", "ok": bool(ok_all), "plane": comm.ext_last_reduce_stats(1)["plane"]})
    elif scenario == "multiple":
        # the per-parameter pattern: many tensors of very different sizes in one call
        sizes = [7, 768, 768 * 4, 50_000, 1, 2_359_296, 3 * 768, 590_000, 13, 768] * 3
        tensors = [make_input(300 + k, rank, n, torch.float32).to(dev) for k, n in enumerate(sizes)]
        descs = []
        for k, t in enumerate(tensors):
            d = pccl.ReduceDescriptor(t.numel(), pccl.ReduceOp.AVG, k, pccl.ReduceOperandDescriptor(pccl.DataType.FLOAT))
            descs.append(pccl.ReduceOpDescriptor.from_torch(t, t, d))
        t0 = time.perf_counter()
        info = comm.all_reduce_multiple_with_retry(descs, max_in_flight=8)
        if dev.type == "cuda":
            torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        ok_all = info.world_size == world
        for k, (t, n) in enumerate(zip(tensors, sizes)):
            want = expected(300 + k, world, n, torch.float32, pccl.ReduceOp.AVG)
            ok_all = ok_all and torch.allclose(t.cpu(), want, rtol=1e-5, atol=1e-5)
        results["cases"].append({"id": "multiple", "ok": bool(ok_all), "ms": round(dt * 1e3, 3), "tensors": len(sizes),
                                 "plane": comm.ext_last_reduce_stats(0)["plane"]})
    elif scenario == "lease":
        # After the first (voted) operation on a tag the group runs that tag without the master.  Checks: results stay
        # right, the stats say "leased", symmetric and in-place operands of awkward sizes, many tags in flight at once,
        # and a host-memory operation on a leased tag (the lease of that tag ends, the operation takes the TCP ring,
        # and the next device operation is voted again).
        import os as _os
        lease_on = _os.environ.get("PCCL_LEASE", "1") != "0"

        def check(name, got, want, st, want_leased):
            nonlocal ok_all
            if want_leased is not None:
                want_leased = want_leased and lease_on
            ok = torch.allclose(got.cpu().double(), want.double(), rtol=1e-5, atol=1e-5) and (want_leased is None or st["leased"] == want_leased)
            ok_all = ok_all and ok
            results["cases"].append({"id": name, "ok": bool(ok), "plane": st["plane"], "leased": st["leased"], "us": round(st["device_ms"] * 1e3, 1)})

        for it, n in enumerate([256, 256, 1000, 70_001, 300_000, 2_000_003, 256]):
            x = make_input(600 + it, rank, n, torch.float32).to(dev)
            y = torch.empty_like(x)
            comm.all_reduce(x, y, op=pccl.ReduceOp.SUM, tag=0)
            check(f"tag0_{it}_{n}", y, expected(600 + it, world, n, torch.float32, pccl.ReduceOp.SUM), comm.ext_last_reduce_stats(0), it > 0)
            comm.all_reduce(x, x, op=pccl.ReduceOp.AVG, tag=0)  # in place on an ordinary buffer
            check(f"tag0_inplace_{it}", x, expected(600 + it, world, n, torch.float32, pccl.ReduceOp.AVG), comm.ext_last_reduce_stats(0), True)
        # symmetric operands, in place: small (one-shot), around the switch-over sizes, counts that are no multiple of four
        for it, n in enumerate([5, 4099, 65_536, 65_541, 262_147, 300_001]):
            sym = comm.ext_symmetric_empty(n, torch.float32)
            for rep in range(3):
                sym.copy_(make_input(700 + it, rank, n, torch.float32))
                comm.all_reduce(sym, sym, op=pccl.ReduceOp.SUM, tag=1)
                check(f"sym_{n}_{rep}", sym, expected(700 + it, world, n, torch.float32, pccl.ReduceOp.SUM), comm.ext_last_reduce_stats(1), None)
            comm.ext_symmetric_free(sym)
        # a view into the middle of a symmetric tensor is not 16-byte aligned: it must be staged, not faulted on
        sym = comm.ext_symmetric_empty(1001, torch.float32)
        sym.copy_(make_input(720, rank, 1001, torch.float32))
        view = sym[1:]
        comm.all_reduce(view, view, op=pccl.ReduceOp.SUM, tag=1)
        check("sym_view", view, expected(720, world, 1001, torch.float32, pccl.ReduceOp.SUM)[1:], comm.ext_last_reduce_stats(1), None)
        ok_all = ok_all and abs(float(sym[0]) - float(make_input(720, rank, 1001, torch.float32)[0])) < 1e-6
        # 16 tags in flight, twice (the second round is leased on every tag)
        import time as _t
        for rnd in range(3):
            xs = [make_input(800 + t, rank, 256, torch.float32).to(dev) for t in range(16)]
            ys = [torch.empty_like(x) for x in xs]
            torch.cuda.synchronize()
            t0 = _t.perf_counter()
            hs = [comm.all_reduce_async(xs[t], ys[t], op=pccl.ReduceOp.SUM, tag=100 + t) for t in range(16)]
            oks = [h.wait()[0] for h in hs]
            dt = _t.perf_counter() - t0
            ok = all(oks) and all(torch.allclose(ys[t].cpu(), expected(800 + t, world, 256, torch.float32, pccl.ReduceOp.SUM), rtol=1e-5, atol=1e-5) for t in range(16))
            ok_all = ok_all and ok
            results["cases"].append({"id": f"16tags_round{rnd}", "ok": bool(ok), "total_us": round(dt * 1e6, 1),
                                     "leased": comm.ext_last_reduce_stats(100)["leased"]})
        # latency of a leased 1 KB operation (host wall clock around the blocking call)
        x = make_input(900, rank, 256, torch.float32).to(dev)
        y = torch.empty_like(x)
        for _ in range(20):
            comm.all_reduce(x, y, op=pccl.ReduceOp.SUM, tag=0)
        torch.cuda.synchronize()
        t0 = _t.perf_counter()
        for _ in range(200):
            comm.all_reduce(x, y, op=pccl.ReduceOp.SUM, tag=0)
        results["leased_1KB_us"] = round((_t.perf_counter() - t0) / 200 * 1e6, 2)
        # enqueued (stream-ordered) operations: a chain on one tag whose every link depends on the one before
        z = torch.ones(100_000, device=dev)
        for _ in range(12):
            comm.all_reduce(z, z, op=pccl.ReduceOp.SUM, tag=0, enqueue=True)
            z.mul_(0.5)
        comm.ext_wait_stream_ops()
        ok = torch.allclose(z.cpu(), torch.full((100_000,), (world * 0.5) ** 12), rtol=1e-5)
        ok_all = ok_all and ok
        results["cases"].append({"id": "enqueued_chain", "ok": bool(ok), "leased": comm.ext_last_reduce_stats(0)["leased"]})
        # host memory on a leased tag: ends the lease of the tag, runs on the TCP ring, then the device takes over again
        h = make_input(910, rank, 5000, torch.float32)
        comm.all_reduce(h, h, op=pccl.ReduceOp.SUM, tag=0)
        st = comm.ext_last_reduce_stats(0)
        ok = torch.allclose(h, expected(910, world, 5000, torch.float32, pccl.ReduceOp.SUM), rtol=1e-5, atol=1e-5) and st["plane"] == "tcp_ring"
        ok_all = ok_all and ok
        results["cases"].append({"id": "host_on_leased_tag", "ok": bool(ok), "plane": st["plane"]})
        for it in range(3):
            comm.all_reduce(x, y, op=pccl.ReduceOp.SUM, tag=0)
            check(f"after_host_{it}", y, expected(900, world, 256, torch.float32, pccl.ReduceOp.SUM), comm.ext_last_reduce_stats(0), it > 0)
    elif scenario == "trusted_hashes":
        n = 3_000_001
        comm.ext_trust_content_hashes(True)
        start = make_input(500, 0, n, torch.float32)
        params = start.clone().to(dev)
        inner = (start - 0.01 * make_input(520, rank, n, torch.float32)).to(dev)
        mom = torch.zeros(n, device=dev)
        state = pccl.SharedState([pccl.TensorInfo.from_torch(params, "params"), pccl.TensorInfo.from_torch(mom, "momentum")])
        state.revision = 1
        info0 = comm.sync_shared_state(state)  # first check: hashes computed by the (batched) hash kernels
        skipped0 = comm.ext_hash_kernels_skipped()
        for k in range(2):
            comm.ext_diloco_outer_step(params, inner, mom, lr=0.7, mu=0.9, first=k == 0, tag=3)
            st = comm.ext_last_reduce_stats(3)
            state.push_revision()
            before = comm.ext_hash_kernels_skipped()
            info = comm.sync_shared_state(state)
            ok = (st["content_hash"] is not None and st["content_hash"][0] == ops.mset64(params) and st["content_hash"][1] == ops.mset64(mom)
                  and comm.ext_hash_kernels_skipped() - before == 2 and info.tx_bytes == 0 and info.rx_bytes == 0 and st["kernel_launches"] == 1)
            ok_all = ok_all and ok
            results["cases"].append({"id": f"trusted_{k}", "ok": bool(ok), "skipped": comm.ext_hash_kernels_skipped() - before, "plane": st["plane"],
                                     "leased": st["leased"]})
            inner.sub_(0.001 * make_input(530 + k, rank, n, torch.float32).to(dev))
        # a tensor the caller changed through the library's back is the caller's business; one the LIBRARY changed loses its record
        comm.all_reduce(mom, mom, op=pccl.ReduceOp.AVG, tag=4)
        before = comm.ext_hash_kernels_skipped()
        state.push_revision()
        comm.sync_shared_state(state)
        ok = comm.ext_hash_kernels_skipped() - before == 1  # params still known, momentum hashed afresh
        ok_all = ok_all and ok and skipped0 == 0 and info0.rx_bytes == 0
        results["cases"].append({"id": "record_dropped_on_library_write", "ok": bool(ok)})
    elif scenario == "overlap":
        # A local compute kernel on another stream runs WHILE a large reduction is in flight: the reduction is stream
        # ordered (no host or device synchronisation) and, with an SM budget below the whole GPU (PCCL_DEVICE_MAX_BLOCKS),
        # leaves SMs to the compute kernel.  Wall time of both together must be clearly below the sum.
        import time as _t
        big = make_input(930, rank, 64_000_000, torch.float32).to(dev)
        a = torch.randn(2048, 2048, device=dev)
        side = torch.cuda.Stream(device=dev)
        reps = [4]

        def burn():
            with torch.cuda.stream(side):
                for _ in range(reps[0]):
                    torch.mm(a, a)

        def timed_ms(fn):
            torch.cuda.synchronize()
            t0 = _t.perf_counter()
            fn()
            torch.cuda.synchronize()
            return (_t.perf_counter() - t0) * 1e3

        for _ in range(3):
            comm.all_reduce(big, big, op=pccl.ReduceOp.AVG, tag=2)
        t_red = min(timed_ms(lambda: comm.all_reduce(big, big, op=pccl.ReduceOp.AVG, tag=2, enqueue=True)) for _ in range(3))
        comm.ext_wait_stream_ops()
        burn()
        t_one = min(timed_ms(burn) for _ in range(3)) / reps[0]
        reps[0] = max(2, int(round(1.5 * t_red / max(t_one, 1e-3))))  # compute about 1.5x as long as the reduction
        t_burn = min(timed_ms(burn) for _ in range(3))

        def both():
            burn()
            comm.all_reduce(big, big, op=pccl.ReduceOp.AVG, tag=2, enqueue=True)

        t_both = min(timed_ms(both) for _ in range(3))
        comm.ext_wait_stream_ops()
        results["overlap_ms"] = {"compute": round(t_burn, 3), "reduce": round(t_red, 3), "both": round(t_both, 3)}
        ok_all = t_both < 0.85 * (t_burn + t_red) and comm.ext_last_reduce_stats(2)["leased"]
        results["cases"].append({"id": "overlap", "ok": bool(ok_all)})
    elif scenario == "kill":
        # the last rank dies without saying goodbye while reductions are running; the others must
        # notice quickly (master abort -> abort word -> kernel exit), keep their input, and go on
        import os
        n = 64_000_000
        x = make_input(400, rank, n, torch.float32).to(dev)
        keep = x.clone()
        done, failures, t_fail = 0, 0, None
        for it in range(400):
            if rank == world - 1 and it == 5:
                os._exit(0)
            try:
                t0 = time.perf_counter()
                comm.all_reduce(x, x, op=pccl.ReduceOp.SUM, tag=0)
                done += 1
                x.copy_(keep)
            except pccl.PCCLError as e:
                failures += 1
                t_fail = time.perf_counter() - t0
                if e.result == pccl.Result.TOO_FEW_PEERS:
                    break
                assert e.result == pccl.Result.RANK_CONNECTION_LOST, e
                assert torch.equal(x, keep), "a failed in-place reduction must leave the input intact"
            if comm.get_attribute(pccl.Attribute.PEER_GROUP_WORLD_SIZE) == world - 1 and done > 20:
                break
        ok_all = failures >= 1 and comm.get_attribute(pccl.Attribute.PEER_GROUP_WORLD_SIZE) == world - 1
        results["cases"].append({"id": "kill", "ok": bool(ok_all), "done": done, "failures": failures,
                                 "seconds_to_fail": None if t_fail is None else round(t_fail, 3)})
    elif scenario == "shared_state":
        # rank 0 holds the truth; the others start from garbage and must end up identical
        n = 5_000_003
        w = make_input(201, 0 if rank == 0 else 7 + rank, n, torch.bfloat16).to(dev)
        m = make_input(202, 0 if rank == 0 else 9 + rank, 1000, torch.float32).to(dev)
        state = pccl.SharedState([pccl.TensorInfo.from_torch(w, "weights"), pccl.TensorInfo.from_torch(m, "momentum")])
        state.revision = 1
        strategy = pccl.SharedStateSyncStrategy.SEND_ONLY if rank == 0 else pccl.SharedStateSyncStrategy.RECEIVE_ONLY
        info = comm.sync_shared_state(state, strategy)
        want_w = make_input(201, 0, n, torch.bfloat16)
        ok_all = torch.equal(w.cpu(), want_w) and torch.equal(m.cpu(), make_input(202, 0, 1000, torch.float32))
        if rank == 0:
            ok_all = ok_all and info.rx_bytes == 0 and info.tx_bytes == (world - 1) * (n * 2 + 4000)
        else:
            ok_all = ok_all and info.rx_bytes == n * 2 + 4000
        # second sync: identical content, nothing moves
        state.revision += 1
        info2 = comm.sync_shared_state(state, pccl.SharedStateSyncStrategy.ENFORCE_POPULAR)
        ok_all = ok_all and info2.tx_bytes == 0 and info2.rx_bytes == 0
        results["cases"].append({"id": "shared_state", "ok": bool(ok_all), "tx": info.tx_bytes, "rx": info.rx_bytes})
    elif scenario == "outer_step":
        # DiLoCo outer step as one collective: every peer starts from the same parameters, drifts differently, and
        # must end with the parameters a direct evaluation of the update gives (two steps: first and steady state)
        n = 5_000_003
        start = make_input(500, 0, n, torch.float32)
        params = start.clone().to(dev)
        inner = torch.empty_like(params)
        mom = torch.full_like(params, 77.0)  # garbage: the first step must not read it
        ref_p, ref_m = start.clone(), torch.zeros(n)
        lr, mu = 0.7, 0.9
        ok_all = True
        for k in range(2):
            drifts = [0.01 * make_input(510 + k, r, n, torch.float32) for r in range(world)]
            inner.copy_((ref_p - drifts[rank]).to(dev))
            g = sum((ref_p - (ref_p - d)) for d in drifts) / world
            ref_m = g.clone() if k == 0 else mu * ref_m + g
            ref_p = ref_p - lr * (g + mu * ref_m)
            info = comm.ext_diloco_outer_step(params, inner, mom, lr=lr, mu=mu, first=k == 0, tag=7)
            st = comm.ext_last_reduce_stats(7)
            ok = (info.world_size == world and torch.allclose(params.cpu(), ref_p, rtol=1e-4, atol=1e-5) and
                  torch.allclose(mom.cpu(), ref_m, rtol=1e-4, atol=1e-5) and torch.equal(inner, params))
            ok_all = ok_all and ok
            results["cases"].append({"id": f"outer_step_{k}", "ok": bool(ok), "plane": st["plane"], "ms": round(st["device_ms"], 4),
                                     "launches": st["kernel_launches"], "hash": ops.simplehash(params)})
    results["ok"] = bool(ok_all)
    print(json.dumps(results), flush=True)
    comm.destroy()
    sys.exit(0 if ok_all else 1)


if __name__ == "__main__":
    main()
