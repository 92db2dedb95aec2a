"""Python surface on the CPU (reference: python/tests/unit_tests/pccl_test.py, numpy_only_tests, pytorch_only_tests,
end_to_end/basic_reduce_test)."""
import subprocess
import sys
import threading
from pathlib import Path

import numpy as np
import pytest
import torch

import pccl_b200 as pccl
from test_gpu_multi import run_peers

REPO = Path(__file__).resolve().parent.parent


_MASTERS = []


def start_master():
    m = pccl.MasterNode("127.0.0.1:0")
    m.run()
    _MASTERS.append(m)
    return m, m.port


@pytest.fixture(autouse=True)
def _stop_masters():
    # a running master that loses its last reference waits to be interrupted (reference semantics): a failing test
    # must not leave one behind
    yield
    while _MASTERS:
        try:
            _MASTERS.pop().interrupt()
        except Exception:
            pass


def join(port, world, **kw):
    c = pccl.Communicator(f"127.0.0.1:{port}", **kw)
    c.connect(n_attempts=10)
    while c.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < world:
        c.update_topology()
    return c


def run_threads(world, fn):
    m, port = start_master()
    errors, results = [], {}

    def body(r):
        try:
            c = join(port, world)
            results[r] = fn(r, c)
            c.destroy()
        except Exception as e:  # noqa: BLE001
            errors.append((r, repr(e)))

    ts = [threading.Thread(target=body, args=(r,)) for r in range(world)]
    [t.start() for t in ts]
    [t.join(timeout=120) for t in ts]
    m.interrupt()
    assert not errors, errors
    return results


def test_import_surface():
    for name in ["Communicator", "MasterNode", "SharedState", "TensorInfo", "ReduceOp", "DataType", "QuantizationOptions",
                 "ReduceDescriptor", "ReduceOpDescriptor", "AsyncReduceHandle", "PCCLError", "Attribute", "SharedStateSyncStrategy",
                 "DistributionHint", "QuantizationAlgorithm", "ReduceOperandDescriptor", "ReduceInfo", "SharedStateSyncInfo"]:
        assert hasattr(pccl, name), name
    assert isinstance(pccl.cuda.is_available(), bool)
    assert pccl.DataType.from_torch_dtype(torch.bfloat16) == pccl.DataType.BFLOAT16
    assert pccl.DataType.from_numpy_dtype(np.float32) == pccl.DataType.FLOAT
    assert pccl.DataType.FLOAT.to_torch_dtype() == torch.float32
    assert pccl.DataType.BFLOAT16.itemsize == 2


def test_master_lifecycle():
    m, port = start_master()
    assert port > 0
    m.interrupt()
    del m
    m2 = pccl.MasterNode("127.0.0.1:0")  # never run: must still be destroyable
    del m2


def test_communicator_lifecycle_and_errors():
    c = pccl.Communicator("127.0.0.1:1")
    with pytest.raises(pccl.PCCLError) as e:
        c.update_topology()  # before connect
    assert e.value.result == pccl.Result.INVALID_USAGE
    with pytest.raises(pccl.PCCLError) as e:
        c.connect(n_attempts=1)  # nobody listens there
    assert e.value.result == pccl.Result.MASTER_CONNECTION_FAILED
    c.destroy()
    c.destroy()  # idempotent
    m, port = start_master()
    solo = pccl.Communicator(f"127.0.0.1:{port}")
    solo.connect()
    assert solo.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) == 1
    x = np.ones(8, dtype=np.float32)
    with pytest.raises(pccl.PCCLError) as e:
        solo.all_reduce(x, x, op=pccl.ReduceOp.SUM)
    assert e.value.result == pccl.Result.TOO_FEW_PEERS
    with pytest.raises(ValueError):
        solo.all_reduce([1, 2], [3, 4], op=pccl.ReduceOp.SUM)
    with pytest.raises(ValueError):
        pccl.TensorInfo.from_torch(torch.ones(4, 4).t(), "non_contiguous")
    solo.destroy()
    m.interrupt()


@pytest.mark.parametrize("world", [2, 3])
def test_numpy_and_torch_all_reduce(world):
    def body(r, c):
        a = np.full(1000, r + 1, dtype=np.float32)
        out = np.zeros_like(a)
        info = c.all_reduce(a, out, op=pccl.ReduceOp.SUM, tag=0)
        assert info.world_size == world and info.tx_bytes > 0
        t = torch.full((333,), float(r), dtype=torch.bfloat16)
        c.all_reduce(t, t, op=pccl.ReduceOp.MAX, tag=1)
        i = torch.arange(10, dtype=torch.int64) * (r + 1)
        c.all_reduce(i, i, op=pccl.ReduceOp.AVG, tag=2)
        h = c.all_reduce_async(a, out, op=pccl.ReduceOp.MIN, tag=3)
        ok, status, _ = h.wait()
        assert ok and status == 0 and h.wait()[0]  # wait() is repeatable
        return out.copy(), t.float().numpy().copy(), i.numpy().copy(), c.ext_last_reduce_stats(0)["plane"]

    res = run_threads(world, body)
    total = sum(range(1, world + 1))
    for r in range(world):
        out, t, i, plane = res[r]
        assert plane == "tcp_ring"
        assert (out == 1).all()  # MIN of (r+1) over ranks
        assert (t == world - 1).all()
        assert (i == (np.arange(10) * total) // world).all()


def test_quantised_uint8_minmax_1000_revisions_style():
    """the reference's python e2e: 2 peers, 1024-element uint8 min-max quantised reductions in a loop"""
    def body(r, c):
        q = pccl.QuantizationOptions(pccl.DataType.UINT8, pccl.QuantizationAlgorithm.MIN_MAX)
        outs = []
        for it in range(30):
            x = torch.linspace(-1, 1, 1024) * (it + 1) + r
            y = torch.empty_like(x)
            c.all_reduce(x, y, op=pccl.ReduceOp.SUM, tag=0, quantization_options=q)
            want = torch.linspace(-1, 1, 1024) * (it + 1) * 2 + 1
            assert (y - want).abs().max() < 0.05 * (it + 1) + 0.05
            outs.append(y.clone())
        return torch.stack(outs)

    res = run_threads(2, body)
    assert torch.equal(res[0], res[1])  # bit-identical on both peers


def test_all_reduce_multiple_with_retry():
    def body(r, c):
        tensors = [torch.full((n,), float(r + 1)) for n in (10, 1000, 100_000, 7)]
        descs = []
        for k, t in enumerate(tensors):
            d = pccl.ReduceDescriptor(t.numel(), pccl.ReduceOp.SUM, k, pccl.ReduceOperandDescriptor(pccl.DataType.FLOAT))
            descs.append(pccl.ReduceOpDescriptor.from_torch(t, t, d))
        info = c.all_reduce_multiple_with_retry(descs, max_in_flight=3)
        assert info.world_size == 3
        return [float(t[0]) for t in tensors]

    for r, vals in run_threads(3, body).items():
        assert vals == [6.0] * 4


def test_shared_state_revisions_and_strategies():
    def body(r, c):
        w = torch.full((5000,), 1.0 if r == 0 else 0.0)
        step = torch.tensor([0], dtype=torch.int64)
        state = pccl.SharedState([pccl.TensorInfo.from_torch(w, "w"), pccl.TensorInfo.from_torch(step, "step")])
        state.revision = 5
        info = c.sync_shared_state(state, pccl.SharedStateSyncStrategy.SEND_ONLY if r == 0 else pccl.SharedStateSyncStrategy.RECEIVE_ONLY)
        first = (info.tx_bytes, info.rx_bytes)
        assert (w == 1.0).all()
        for it in range(3):
            w += 1
            step += 1
            state.push_revision()
            info = c.sync_shared_state(state)
            assert info.tx_bytes == 0 and info.rx_bytes == 0
        return first, state.revision

    res = run_threads(2, body)
    assert res[0][0] == (20000, 0) and res[1][0] == (0, 20000)
    assert res[0][1] == res[1][1] == 8


@pytest.mark.parametrize("scenario", ["allreduce", "concurrent", "shared_state", "multiple", "outer_step"])
def test_multi_process_peers_on_cpu(scenario):
    for r in run_peers(3, scenario, devices=["cpu"] * 3):
        assert r["ok"], r


def test_two_level_ring_between_pretended_hosts():
    """PCCL_HOST_ID tells peers which machine they are on; two "hosts" of two peers reduce over a local ring and a ring
    between hosts; quantised operations too (codes travel between the hosts only)."""
    hosts = [{"PCCL_HOST_ID": h} for h in ("left", "right", "left", "right")]
    results = run_peers(4, "allreduce", devices=["cpu"] * 4, rank_env=hosts)
    by_case = {}
    planes = set()
    for r in results:
        assert r["ok"], r
        for c in r["cases"]:
            by_case.setdefault(c["id"], set()).add(c["hash"])
            planes.add(c["plane"])
    assert planes == {"tcp_hierarchical"}, planes
    assert all(len(h) == 1 for h in by_case.values()), "peers disagree bit-wise"
    # the DiLoCo outer step of host tensors reduces its pseudo-gradients the same way
    stepped = run_peers(4, "outer_step", devices=["cpu"] * 4, rank_env=hosts)
    assert all(r["ok"] for r in stepped), stepped
    assert {c["plane"] for r in stepped for c in r["cases"]} == {"tcp_hierarchical"}
    assert len({c["hash"] for r in stepped for c in r["cases"] if c["id"] == "outer_step_1"}) == 1
    # the same peers without host ids on loopback: one flat ring
    flat = run_peers(4, "concurrent", devices=["cpu"] * 4, env={"PCCL_HIERARCHY": "0"}, rank_env=hosts)
    assert all(r["ok"] for r in flat) and {c["plane"] for r in flat for c in r["cases"]} == {"tcp_ring"}
    # ... and two operations in flight at once share the two extra links of the hierarchy
    both = run_peers(4, "concurrent", devices=["cpu"] * 4, rank_env=hosts)
    assert all(r["ok"] for r in both) and {c["plane"] for r in both for c in r["cases"]} == {"tcp_hierarchical"}


@pytest.mark.parametrize("hash_type", ["mset", "crc32"])
def test_shared_state_with_the_other_hash_types(hash_type):
    """PCCL_SHARED_STATE_HASH switches the content hash of the whole run: mset64 (the order-independent hash the collective
    kernels produce as a by-product; host twin here) and CRC32C decide "who is outdated" exactly like simplehash does."""
    for r in run_peers(3, "shared_state", devices=["cpu"] * 3, env={"PCCL_SHARED_STATE_HASH": hash_type}):
        assert r["ok"], r


def test_stream_api_degrades_gracefully_on_host_tensors():
    """all_reduce(enqueue=True) on host tensors is the ordinary blocking call; waiting for enqueued operations when there are
    none succeeds."""
    def body(r, c):
        x = torch.full((1000,), float(r + 1))
        c.all_reduce(x, x, op=pccl.ReduceOp.SUM, tag=3, enqueue=True)
        c.ext_wait_stream_ops()
        c.ext_wait_stream_ops(3)
        c.ext_trust_content_hashes(True)
        assert c.ext_hash_kernels_skipped() == 0
        st = c.ext_last_reduce_stats(3)
        assert st["plane"] == "tcp_ring" and not st["leased"] and st["content_hash"] is None
        return float(x[0])

    assert run_threads(2, body) == {0: 3.0, 1: 3.0}


def test_diloco_cpu_matches_torch_sgd():
    from pccl_b200.parallel import DiLoCo

    def body(r, c):
        torch.manual_seed(0)
        params = torch.randn(2000)
        ref = params.clone().requires_grad_(True)
        opt = torch.optim.SGD([ref], lr=0.7, momentum=0.9, nesterov=True)
        d = DiLoCo(c, params, lr=0.7, momentum=0.9)
        for step in range(3):
            inner = params - 0.01 * (step + 1) * (r + 1)   # what this peer's inner optimiser produced
            d.outer_step(inner)
            # plain PyTorch reference: averaged pseudo-gradient = 0.01*(step+1)*mean(r+1) = 0.015*(step+1) for two peers
            ref.grad = torch.full_like(ref, 0.015 * (step + 1))
            opt.step()
            torch.testing.assert_close(params, ref.detach(), rtol=1e-5, atol=1e-6)
            assert torch.equal(inner, params)  # the inner copy restarts from the new outer parameters
        return True

    assert all(run_threads(2, body).values())


def test_peer_killed_mid_run_recovers_quickly():
    """a peer exits without notice while reductions run (the reference's stress orchestrator SIGKILLs peers)"""
    for r in run_peers(3, "kill", devices=["cpu"] * 3, timeout=150):
        c = r["cases"][0]
        assert c["ok"], c
        assert c["seconds_to_fail"] < 5, c


def _ipv6_available() -> bool:
    import socket
    try:
        s = socket.socket(socket.AF_INET6, socket.SOCK_STREAM)
        s.bind(("::1", 0))
        s.close()
        return True
    except OSError:
        return False


@pytest.mark.skipif(not _ipv6_available(), reason="no IPv6 loopback")
def test_everything_over_ipv6_loopback():
    """Master, p2p ring, shared-state server: all reachable through [::1] when the master address is IPv6."""
    with pccl.MasterNode("[::1]:0") as m:
        m.run()
        port = m.port
        errors, results = [], {}

        def body(rank):
            try:
                c = pccl.Communicator(f"[::1]:{port}", p2p_listen_port=48940 + 8 * rank, shared_state_listen_port=48941 + 8 * rank,
                                      benchmark_listen_port=48942 + 8 * rank)
                c.connect(n_attempts=10)
                while c.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < 2:
                    c.update_topology()
                x = torch.full((50_000,), float(rank + 1))
                y = torch.empty_like(x)
                c.all_reduce(x, y, op=pccl.ReduceOp.SUM)
                w = torch.full((10_000,), 7.0 if rank == 0 else -1.0)
                st = pccl.SharedState([pccl.TensorInfo.from_torch(w, "w")])
                st.revision = 1
                info = c.sync_shared_state(st, pccl.SharedStateSyncStrategy.SEND_ONLY if rank == 0 else pccl.SharedStateSyncStrategy.RECEIVE_ONLY)
                results[rank] = (float(y[0]), float(w[-1]), info.rx_bytes)
                c.destroy()
            except Exception as e:  # noqa: BLE001
                errors.append(e)

        ts = [threading.Thread(target=body, args=(r,)) for r in range(2)]
        for t in ts:
            t.start()
        for t in ts:
            t.join(timeout=120)
    assert not errors, errors
    assert results[0] == (3.0, 7.0, 0) and results[1] == (3.0, 7.0, 40_000), results


def test_async_diloco_applies_each_average_one_step_late():
    """AsyncDiLoCo: the reduction of step t overlaps the next inner phase and is applied at step t+1 (flush applies the last
    one).  With pseudo-gradients that do not depend on the parameters the end result must equal plain Nesterov-SGD over the
    same sequence of averages, and all peers must agree bit for bit."""
    from pccl_b200.parallel import AsyncDiLoCo

    def body(r, c):
        params = torch.randn(3000, generator=torch.Generator().manual_seed(0))  # (the global generator is shared between threads)
        ref = params.clone().requires_grad_(True)
        opt = torch.optim.SGD([ref], lr=0.7, momentum=0.9, nesterov=True)
        d = AsyncDiLoCo(c, params, lr=0.7, momentum=0.9)
        applied_before = []
        for step in range(4):
            inner = params - 0.01 * (step + 1) * (r + 1)
            assert d.outer_step(inner)
            applied_before.append(d.applied)
            assert torch.equal(inner, params)
        assert applied_before == [0, 1, 2, 3]  # step t applies the average of step t-1
        d.flush()
        assert d.applied == 4
        for step in range(4):
            ref.grad = torch.full_like(ref, 0.015 * (step + 1))
            opt.step()
        torch.testing.assert_close(params, ref.detach(), rtol=1e-5, atol=1e-6)
        return params.clone()

    res = run_threads(2, body)
    assert torch.equal(res[0], res[1])


def test_gradient_averager_cpu():
    """GradientAverager: flatten .grad, average over the group, scatter back; parameters without a gradient count as zero."""
    from pccl_b200.parallel import GradientAverager

    def body(r, c):
        torch.manual_seed(1)
        model = torch.nn.Sequential(torch.nn.Linear(8, 16), torch.nn.Linear(16, 4))
        for i, p in enumerate(model.parameters()):
            p.grad = torch.full_like(p, float((r + 1) * (i + 1)))
        list(model.parameters())[-1].grad = None  # e.g. an unused head
        avg = GradientAverager(c, model.parameters())
        assert avg.average()
        params = list(model.parameters())
        for i, p in enumerate(params[:-1]):
            assert torch.allclose(p.grad, torch.full_like(p, 1.5 * (i + 1)))
        assert params[-1].grad is None
        return True

    assert all(run_threads(2, body).values())


def test_package_works_without_torch():
    """numpy-only installation (reference: python/tests/numpy_only_tests/numpy_only_test.py): importing, a master, a
    communicator, the single-peer error and a two-peer numpy reduction, with `torch` made unimportable."""
    code = r'''
import sys, threading
sys.modules["torch"] = None
sys.path.insert(0, sys.argv[1])
import numpy as np
import pccl_b200 as pccl
with pccl.MasterNode("127.0.0.1:0") as m:
    m.run()
    lone = pccl.Communicator(f"127.0.0.1:{m.port}", p2p_listen_port=48960, shared_state_listen_port=48961, benchmark_listen_port=48962)
    lone.connect()
    arr = np.random.rand(27)
    try:
        lone.all_reduce_async(arr, arr, op=pccl.ReduceOp.SUM)
        raise SystemExit("a single peer must not be able to reduce")
    except pccl.PCCLError:
        pass
    out = {}
    def second():
        c = pccl.Communicator(f"127.0.0.1:{m.port}", p2p_listen_port=48970, shared_state_listen_port=48971, benchmark_listen_port=48972)
        c.connect(n_attempts=10)
        x = np.full(1000, 2.0, dtype=np.float32); y = np.empty_like(x)
        c.all_reduce(x, y, op=pccl.ReduceOp.SUM)
        out[1] = float(y[0])
        c.destroy()
    t = threading.Thread(target=second); t.start()
    while lone.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < 2:
        lone.update_topology()
    x = np.full(1000, 1.0, dtype=np.float32); y = np.empty_like(x)
    lone.all_reduce(x, y, op=pccl.ReduceOp.SUM)
    t.join(timeout=60)
    assert float(y[0]) == 3.0 and out[1] == 3.0, (y[0], out)
    lone.destroy()
print("NUMPY ONLY OK")
'''
    proc = subprocess.run([sys.executable, "-c", code, str(REPO)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=180)
    assert proc.returncode == 0 and "NUMPY ONLY OK" in proc.stdout, proc.stdout[-3000:]


def test_package_works_without_numpy():
    """torch-only installation (reference: python/tests/pytorch_only_tests/pytorch_only_test.py): the same with `numpy`
    made unimportable and torch tensors as operands and shared state."""
    code = r'''
import sys, threading, warnings
warnings.filterwarnings("ignore")
sys.modules["numpy"] = None
sys.path.insert(0, sys.argv[1])
import torch
import pccl_b200 as pccl
with pccl.MasterNode("127.0.0.1:0") as m:
    m.run()
    lone = pccl.Communicator(f"127.0.0.1:{m.port}")
    lone.connect()
    arr = torch.rand(27)
    try:
        lone.all_reduce_async(arr, arr, op=pccl.ReduceOp.SUM)
        raise SystemExit("a single peer must not be able to reduce")
    except pccl.PCCLError:
        pass
    out = {}
    def second():
        c = pccl.Communicator(f"127.0.0.1:{m.port}")
        c.connect(n_attempts=10)
        w = torch.zeros(100)
        info = c.sync_shared_state(pccl.SharedState([pccl.TensorInfo.from_torch(w, "w")]))
        x = torch.full((1000,), 2.0)
        c.all_reduce(x, x, op=pccl.ReduceOp.AVG)
        out[1] = (float(x[0]), float(w[7]), info.rx_bytes)
        c.destroy()
    t = threading.Thread(target=second); t.start()
    while lone.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < 2:
        lone.update_topology()
    w = torch.full((100,), 5.0)
    state = pccl.SharedState([pccl.TensorInfo.from_torch(w, "w")])
    state.revision = 1                       # ahead of the newcomer: this content is what the group holds
    lone.sync_shared_state(state)
    x = torch.full((1000,), 1.0)
    lone.all_reduce(x, x, op=pccl.ReduceOp.AVG)
    t.join(timeout=60)
    assert float(x[0]) == 1.5 and out[1] == (1.5, 5.0, 400), (x[0], out)
    lone.destroy()
print("TORCH ONLY OK")
'''
    proc = subprocess.run([sys.executable, "-c", code, str(REPO)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=180)
    assert proc.returncode == 0 and "TORCH ONLY OK" in proc.stdout, proc.stdout[-3000:]
