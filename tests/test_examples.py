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
def test_diloco_example_two_peers_cpu(master, mode):
    procs = [subprocess.Popen([sys.executable, str(REPO / "examples" / "diloco.py"), "--master", f"127.0.0.1:{master.port}", "--device", "cpu",
                               "--outer-steps", "2", "--inner-steps", "2", "--min-peers", "2", "--mode", mode],
                              stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=_env()) for _ in range(2)]
    for p in procs:
        out, _ = p.communicate(timeout=300)
        assert p.returncode == 0, out[-4000:]
        assert "outer step 1" in out and "world 2" in out, out[-2000:]


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
