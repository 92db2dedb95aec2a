"""The example programs run end to end on CPU (tiny model, synthetic tokens), as separate processes against a real master."""
import os
import re
import subprocess
import sys
from pathlib import Path

import pytest

import pccl_b200 as pccl

REPO = Path(__file__).resolve().parent.parent


@pytest.fixture
def master():
    with pccl.MasterNode("127.0.0.1:0") as m:  # leaving the block interrupts the master, also when a check fails
        m.run()
        yield m


def _env():
    env = dict(os.environ)
    env["OMP_NUM_THREADS"] = "2"
    return env


def test_fsdp_islands_share_shards_through_peer_groups(master):
    """Two islands of two FSDP ranks each: equal shard ranks form one PCCL peer group and must end every outer step
    with identical shard contents (reference: python/examples/nanogpt_diloco/sync_diloco_fsdp.py)."""
    procs = []
    for island in range(2):
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
               "--master-port", str(29610 + island), str(REPO / "examples" / "diloco_fsdp.py"), "--pccl-master", f"127.0.0.1:{master.port}",
               "--device", "cpu", "--outer-steps", "2", "--inner-steps", "2", "--min-islands", "2", "--p2p-port-base", str(48300 + 64 * island)]
        procs.append(subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=_env()))
    outs = []
    for p in procs:
        out, _ = p.communicate(timeout=300)
        assert p.returncode == 0, out[-4000:]
        outs.append(out)
    digests = [dict(), dict()]
    for island, out in enumerate(outs):
        for m in re.finditer(r"\[island rank (\d+)\] outer step (\d+): .* islands (\d+) .* digest (\d+)", out):
            assert m.group(3) == "2", out
            digests[island][(int(m.group(1)), int(m.group(2)))] = int(m.group(4))
    assert len(digests[0]) == 4 and digests[0] == digests[1], digests
    # the two shard ranks hold different halves of the model
    assert digests[0][(0, 1)] != digests[0][(1, 1)]


@pytest.mark.parametrize("mode", ["sync", "async"])
def test_diloco_example_two_peers_cpu(master, mode, tmp_path):
    procs = [subprocess.Popen([sys.executable, str(REPO / "examples" / "diloco.py"), "--master", f"127.0.0.1:{master.port}", "--device", "cpu",
                               "--outer-steps", "2", "--inner-steps", "2", "--min-peers", "2", "--mode", mode,
                               "--timeline", str(tmp_path / f"peer{i}.svg")],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=_env()) for i in range(2)]
    for p in procs:
        out, _ = p.communicate(timeout=300)
        assert p.returncode == 0, out[-4000:]
        assert "outer step 1" in out and "world 2" in out, out[-2000:]
    for i in range(2):  # the section timeline and its trace (reference: profiler.export_timeline)
        assert "inner" in (tmp_path / f"peer{i}.svg").read_text()
        assert '"traceEvents"' in (tmp_path / f"peer{i}.trace.json").read_text()


def test_ddp_examples_cpu(master):
    """examples/ddp.py (two fault-tolerant peers) and the NCCL/gloo baseline examples/ddp_nccl.py both train."""
    procs = [subprocess.Popen([sys.executable, str(REPO / "examples" / "ddp.py"), "--master", f"127.0.0.1:{master.port}", "--device", "cpu",
                               "--steps", "3", "--min-peers", "2"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=_env())
             for _ in range(2)]
    for p in procs:
        out, _ = p.communicate(timeout=300)
        assert p.returncode == 0, out[-4000:]
    base = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                           "--master-port", "29640", str(REPO / "examples" / "ddp_nccl.py"), "--device", "cpu", "--steps", "3"],
                          stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=_env(), timeout=300)
    assert base.returncode == 0 and "step 2: loss" in base.stdout, base.stdout[-4000:]


def test_diloco_example_checkpoint_and_resume(master, tmp_path):
    """Config file, evaluation, checkpoint, and a second run that resumes: the resumed peers bring their shared-state
    revision along (the first synchronisation of a group may carry any revision) and continue the step count
    (reference: sync_diloco.py:247-262,445-457)."""
    cfg = tmp_path / "cfg.json"
    cfg.write_text('{"model": "tiny", "inner_steps": 2, "outer_steps": 3, "batch_size": 2, "eval_interval": 1, "eval_iters": 1, '
                   f'"checkpoint_dir": "{tmp_path}/ckpt_RANK"' + "}")

    def run(extra):
        procs = []
        for r in range(2):
            c = tmp_path / f"cfg{r}.json"
            c.write_text(cfg.read_text().replace("ckpt_RANK", f"ckpt_{r}"))
            procs.append(subprocess.Popen([sys.executable, str(REPO / "examples" / "diloco.py"), "--master", f"127.0.0.1:{master.port}", "--device", "cpu",
                                           "--config", str(c), "--min-peers", "2"] + extra, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True,
                                          env=_env()))
        outs = []
        for p in procs:
            out, _ = p.communicate(timeout=300)
            assert p.returncode == 0, out[-4000:]
            outs.append(out)
        return outs

    first = run([])
    assert all("val loss" in o and "run step 3" in o for o in first), first[0][-2000:]
    assert (tmp_path / "ckpt_0" / "ckpt.pt").exists() and (tmp_path / "ckpt_1" / "ckpt.pt").exists()
    second = run(["--resume"])
    for o in second:
        assert "resumed from" in o and "at outer step 3" in o, o[-2000:]
        assert "run step 6" in o and "world 2" in o, o[-2000:]


def _run_with_a_late_joiner(base, tmp_path, timeout):
    """Two peers train; a third is started once the run is demonstrably under way (the first peer has printed its third
    outer step) and does 12 outer steps.  The old peers start counting their 15 steps when they first see three peers
    (--count-from-world), so nothing here depends on how fast this machine trains or starts a Python process.  The old peers
    write to files: nobody reads their pipes while the test waits for the newcomer."""
    import time
    logs = [tmp_path / f"old{i}.log" for i in range(2)]
    old = [subprocess.Popen(base + ["--outer-steps", "15", "--min-peers", "2", "--count-from-world", "3"], stdout=open(log, "w"),
                            stderr=subprocess.STDOUT, env=_env()) for log in logs]
    deadline = time.time() + 120
    while "outer step 2:" not in logs[0].read_text():
        assert old[0].poll() is None and time.time() < deadline, logs[0].read_text()[-3000:]
        time.sleep(0.1)
    new = subprocess.Popen(base + ["--outer-steps", "12", "--min-peers", "3"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=_env())
    try:
        out_new, _ = new.communicate(timeout=timeout)
        for p in old:
            p.wait(timeout=timeout)
    finally:
        for p in old + [new]:
            if p.poll() is None:
                p.kill()
    return new.returncode, out_new, [(p.returncode, log.read_text()) for p, log in zip(old, logs)]


def test_sync_diloco_late_joiner(master, tmp_path):
    """A peer that joins a synchronous run mid-way takes part from the next outer step with the run's parameters."""
    base = [sys.executable, str(REPO / "examples" / "diloco.py"), "--master", f"127.0.0.1:{master.port}", "--device", "cpu", "--mode", "sync",
            "--inner-steps", "2", "--batch", "2"]
    rc, out_new, olds = _run_with_a_late_joiner(base, tmp_path, 300)
    assert rc == 0, out_new[-4000:]
    assert "world 3" in out_new, out_new[-3000:]
    first = re.search(r"outer step 0: .* run step (\d+)", out_new)
    assert first and int(first.group(1)) > 1, out_new[:2000]
    for rc_old, out in olds:
        assert rc_old == 0, out[-4000:]
        assert "world 3" in out, out[-3000:]


def test_async_diloco_late_joiner_is_inserted_into_the_pipeline(master, tmp_path):
    """Two peers train in the one-step-delayed mode; a third joins mid-run.  The members apply the reduction the joiner
    missed and re-send the shared state (SEND_ONLY), the joiner receives it (RECEIVE_ONLY) -- reference:
    python/examples/nanogpt_diloco/async_diloco.py:551-580 -- and all three go on together."""
    base = [sys.executable, str(REPO / "examples" / "diloco.py"), "--master", f"127.0.0.1:{master.port}", "--device", "cpu", "--mode", "async",
            "--inner-steps", "2", "--batch", "2"]
    rc, out_new, olds = _run_with_a_late_joiner(base, tmp_path, 400)
    assert rc == 0, out_new[-4000:]
    assert "world 3" in out_new, out_new[-3000:]
    # the joiner adopted the run's step counter: its first line is not "run step 1"
    first = re.search(r"outer step 0: .* run step (\d+)", out_new)
    assert first and int(first.group(1)) > 1, out_new[:2000]
    for rc_old, out in olds:
        assert rc_old == 0, out[-4000:]
        assert "world 3" in out, out[-3000:]
