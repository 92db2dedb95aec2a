"""Multi-process peers, one per GPU, through the whole stack (master, control plane, NVLink kernels).
Needs at least two GPUs; skipped otherwise."""
import json
import subprocess
import sys
from pathlib import Path

import pytest

import pccl_b200 as pccl
from conftest import gpu_count

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
HERE = Path(__file__).resolve().parent


def run_peers(world: int, scenario: str, devices=None, timeout=600, env=None, rank_env=None):
    """`env` goes to every peer, `rank_env` (a list of dicts) to one peer each."""
    import os
    with pccl.MasterNode("127.0.0.1:0") as master:  # leaving the block interrupts the master, also when a check fails
        master.run()
        port = master.port
        devices = devices or [f"cuda:{i}" for i in range(world)]
        procs = [subprocess.Popen([sys.executable, str(HERE / "peer_worker.py"), str(port), str(r), str(world), devices[r], scenario],
                                  stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                                  env=dict(os.environ, **(env or {}), **(rank_env[r] if rank_env else {})))
                 for r in range(world)]
        outs = []
        for p in procs:
            try:
                o, e = p.communicate(timeout=timeout)
            except subprocess.TimeoutExpired:
                for q in procs:
                    q.kill()
                raise
            outs.append((p.returncode, o, e))
    results = []
    for rank, (rc, o, e) in enumerate(outs):
        if scenario == "kill" and rank == world - 1:
            continue  # the victim exits without a report
        assert rc == 0, f"peer failed rc={rc}\nstdout:{o[-3000:]}\nstderr:{e[-3000:]}"
        results.append(json.loads(o.strip().splitlines()[-1]))
    return results


def worlds():
    n = gpu_count()
    return [w for w in (2, 4, 8) if w <= n]


@pytest.mark.parametrize("world", worlds() or [pytest.param(2, marks=pytest.mark.skip(reason="needs >= 2 GPUs"))])
def test_allreduce_over_nvlink(world):
    results = run_peers(world, "allreduce")
    by_case = {}
    for r in results:
        for c in r["cases"]:
            assert c["ok"], c
            assert c["plane"].startswith("p2p") or c["plane"] == "nvls", f"case {c['id']} ran on {c['plane']}, expected the NVLink plane"
            by_case.setdefault(c["id"], set()).add(c["hash"])
    # every peer must hold bit-identical results
    for cid, hashes in by_case.items():
        assert len(hashes) == 1, f"case {cid}: peers disagree bit-wise"


@pytest.mark.parametrize("world", worlds()[:1] or [pytest.param(2, marks=pytest.mark.skip(reason="needs >= 2 GPUs"))])
def test_concurrent_tags_over_nvlink(world):
    for r in run_peers(world, "concurrent"):
        assert r["ok"]


@pytest.mark.parametrize("world", worlds()[:1] or [pytest.param(2, marks=pytest.mark.skip(reason="needs >= 2 GPUs"))])
def test_many_small_tensors_are_fused(world):
    for r in run_peers(world, "multiple"):
        assert r["ok"], r
        assert r["cases"][0]["plane"].startswith("p2p") or r["cases"][0]["plane"] == "nvls"


@pytest.mark.parametrize("world", worlds()[:2] or [pytest.param(2, marks=pytest.mark.skip(reason="needs >= 2 GPUs"))])
def test_shared_state_device_tensors(world):
    for r in run_peers(world, "shared_state"):
        assert r["ok"], r


@pytest.mark.parametrize("world", [w for w in worlds() if w >= 3][:1] or [pytest.param(3, marks=pytest.mark.skip(reason="needs >= 3 GPUs"))])
def test_peer_killed_mid_reduction_on_the_device_plane(world):
    for r in run_peers(world, "kill", timeout=180):
        c = r["cases"][0]
        assert c["ok"], c
        assert c["seconds_to_fail"] is not None and c["seconds_to_fail"] < 10, c


@pytest.mark.parametrize("world", worlds()[:2] or [pytest.param(2, marks=pytest.mark.skip(reason="needs >= 2 GPUs"))])
def test_fused_diloco_outer_step(world):
    """pseudo-gradient -> average -> Nesterov update as ONE kernel per peer (P2P two-shot at 2 GPUs, NVLS from 4)."""
    results = run_peers(world, "outer_step")
    for r in results:
        assert r["ok"], r
        for c in r["cases"]:
            assert c["launches"] == 1 and (c["plane"] == "nvls" if world >= 4 else c["plane"] == "p2p_two_shot"), c
    assert len({c["hash"] for r in results for c in r["cases"] if c["id"] == "outer_step_1"}) == 1, "peers must agree bit for bit"


@pytest.mark.parametrize("world", worlds()[:1] or [pytest.param(2, marks=pytest.mark.skip(reason="needs >= 2 GPUs"))])
def test_outer_step_larger_than_a_staging_slot(world):
    """More elements than one staging slot holds: the library falls back to scratch buffer + plain reduction (several
    launches on the device plane) + update, with the same result."""
    results = run_peers(world, "outer_step", env={"PCCL_ARENA_SLOT_MB": "8", "PCCL_SYMMETRIC_MB": "0"})
    for r in results:
        assert r["ok"], r
        assert all(c["launches"] > 1 for c in r["cases"]), r
    assert len({c["hash"] for r in results for c in r["cases"] if c["id"] == "outer_step_1"}) == 1


@pytest.mark.parametrize("world", [w for w in worlds() if w >= 4][:1] or [pytest.param(4, marks=pytest.mark.skip(reason="needs >= 4 GPUs (NVLS)"))])
def test_pipelined_nvls_kernel(world):
    """The warp-specialised pipeline (copy-in || switch reduction || copy-out) on every staged NVLS operation, including the
    DiLoCo outer step, with the same expectations as the plain kernels."""
    env = {"PCCL_KERNEL_PIPELINE": "1", "PCCL_PIPELINE_MIN_MB": "0"}
    for r in run_peers(world, "allreduce", env=env):
        assert r["ok"], r
        assert any(c["plane"] == "nvls" for c in r["cases"]), r
    results = run_peers(world, "outer_step", env=env)
    for r in results:
        assert r["ok"], r
        assert all(c["launches"] == 1 and c["plane"] == "nvls" for c in r["cases"]), r
    assert len({c["hash"] for r in results for c in r["cases"] if c["id"] == "outer_step_1"}) == 1


@pytest.mark.parametrize("world", worlds()[:2] or [pytest.param(2, marks=pytest.mark.skip(reason="needs >= 2 GPUs"))])
def test_leased_operations(world):
    """Operations after the first one on a tag run without the master (lease + same-box sequencer)."""
    results = run_peers(world, "lease", timeout=300)
    for r in results:
        assert r["ok"], [c for c in r["cases"] if not c["ok"]]
    print({"world": world, "leased_1KB_us": [r["leased_1KB_us"] for r in results],
           "16tags_us": [c["total_us"] for c in results[0]["cases"] if str(c["id"]).startswith("16tags")],
           })


@pytest.mark.parametrize("world", worlds()[:1] or [pytest.param(2, marks=pytest.mark.skip(reason="needs >= 2 GPUs"))])
def test_everything_still_works_without_leases(world):
    for r in run_peers(world, "lease", timeout=300, env={"PCCL_LEASE": "0"}):
        assert r["ok"], [c for c in r["cases"] if not c["ok"]]
        assert not any(c.get("leased") for c in r["cases"])


@pytest.mark.parametrize("world", worlds()[:1] or [pytest.param(2, marks=pytest.mark.skip(reason="needs >= 2 GPUs"))])
def test_reduction_overlaps_with_compute_on_another_stream(world):
    results = run_peers(world, "overlap", timeout=300, env={"PCCL_DEVICE_MAX_BLOCKS": "32"})
    print({"overlap_ms": [r.get("overlap_ms") for r in results]})
    for r in results:
        assert r["ok"], r


@pytest.mark.parametrize("mode", ["sync", "async"])
def test_diloco_training_with_a_late_joiner_on_gpus(mode, tmp_path):
    """Training-level check on the device plane (the reference's MNIST tier with a late joiner and CUDA shared state,
    python/tests/end_to_end/mnist_ddp/mnist_ddp_e2e_test.py): two peers train a small GPT with DiLoCo on two GPUs, a third
    process joins mid-run, everybody finishes, the loss goes down and the joiner took over the run's step counter."""
    import os
    import re
    import time
    if gpu_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    n = gpu_count()
    repo = HERE.parent
    logs = [tmp_path / f"peer{i}.log" for i in range(3)]
    with pccl.MasterNode("127.0.0.1:0") as master:
        master.run()
        base = [sys.executable, "-u", str(repo / "examples" / "diloco.py"), "--master", f"127.0.0.1:{master.port}", "--mode", mode, "--inner-steps", "4",
                "--batch", "4"]
        env = dict(os.environ, PCCL_DEVICE_TIMEOUT_MS="8000", PCCL_LOG_LEVEL="INFO")

        def start(i, device, steps, min_peers):
            return subprocess.Popen(base + ["--device", device, "--outer-steps", str(steps), "--min-peers", str(min_peers)], stdout=open(logs[i], "w"),
                                    stderr=subprocess.STDOUT, env=env)

        # on a B200 an outer step of the small model takes ~20 ms: the old peers must still be at it when the newcomer arrives
        procs = [start(0, "cuda:0", 1500, 2), start(1, "cuda:1", 1500, 2)]
        time.sleep(3)
        procs.append(start(2, f"cuda:{2 if n > 2 else 1}", 150, 3))
        deadline = time.time() + 150
        while time.time() < deadline and any(p.poll() is None for p in procs):
            time.sleep(0.5)
        hung = [i for i, p in enumerate(procs) if p.poll() is None]
        for p in procs:
            if p.poll() is None:
                p.kill()
    outs = [log.read_text() for log in logs]
    tails = "\n".join(f"--- peer {i} (rc {procs[i].returncode}) ---\n{o[-1800:]}" for i, o in enumerate(outs))
    assert not hung, f"peers {hung} did not finish\n{tails}"
    assert all(p.returncode == 0 for p in procs), tails
    out_new = outs[2]
    assert "world 3" in out_new, tails
    first = re.search(r"outer step 0: .* run step (\d+)", out_new)
    assert first and int(first.group(1)) > 1, out_new[:1500]
    for out in outs[:2]:
        losses = [float(m) for m in re.findall(r"outer step \d+: loss ([0-9.]+)", out)]
        assert len(losses) == 1500 and sum(losses[-5:]) / 5 < sum(losses[:5]) / 5 - 0.5, losses[:5] + losses[-5:]
        assert "world 3" in out
        if n > 2 and mode == "sync":
            assert "plane p2p" in out or "plane nvls" in out, "peers on distinct GPUs must reduce on the device plane"


@pytest.mark.parametrize("world", worlds()[:2] or [pytest.param(2, marks=pytest.mark.skip(reason="needs >= 2 GPUs"))])
def test_outer_step_hashes_make_the_next_shared_state_check_free(world):
    """With the mset64 hash type and trusted content hashes the fused outer step hashes parameters and momentum while it
    writes them; the following shared-state synchronisation launches no hash kernel for them and moves nothing."""
    for r in run_peers(world, "trusted_hashes", timeout=200, env={"PCCL_SHARED_STATE_HASH": "mset"}):
        assert r["ok"], r


@pytest.mark.skipif(gpu_count() < 2, reason="needs >= 2 GPUs")
def test_cuda_example_program_on_two_gpus(tmp_path):
    """examples/cuda/stream_allreduce.cu, built with nvcc and run as one process per GPU: symmetric buffers, stream-ordered
    enqueued all-reduces between two kernels, no host synchronisation inside the loop.  The weights it prints follow from the
    arithmetic of its two kernels."""
    import re
    from test_native_examples import build_cuda_example
    exe = build_cuda_example(tmp_path)
    steps = 12
    with pccl.MasterNode("127.0.0.1:0") as master:
        master.run()
        procs = [subprocess.Popen([str(exe), f"127.0.0.1:{master.port}", str(g), str(steps), "2"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
                 for g in range(2)]
        outs = []
        for p in procs:
            try:
                outs.append(p.communicate(timeout=150)[0])
            except subprocess.TimeoutExpired:
                for q in procs:
                    q.kill()
                raise
    assert all(p.returncode == 0 for p in procs), outs
    want = [-0.01 * sum((i + s) % 5 + 0.5 for s in range(steps)) for i in range(4)]  # average of the seeds 0 and 1 is 0.5
    for o in outs:
        m = re.search(r"with 2 peers, ([\d.]+) ms per step .* plane (\d), leased (\d), weights\[0..3\] = (.*)", o)
        assert m, o
        assert int(m.group(2)) in (2, 3, 4) and int(m.group(3)) == 1, o  # an NVLink plane, under a lease
        got = [float(x) for x in m.group(4).split()]
        assert all(abs(a - b) < 2e-3 for a, b in zip(got, want)), (got, want)
