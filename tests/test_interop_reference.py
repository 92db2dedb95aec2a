"""Wire compatibility with the unmodified reference build (baseline/_ref, see baseline/README.md): masters and peers of
both implementations are mixed in one run.  Skipped when the reference is not installed."""
import subprocess
import sys
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parent.parent
SCRIPT = REPO / "tests" / "interop_peer.py"
pytestmark = pytest.mark.skipif(not (REPO / "baseline" / "_ref" / "pccl").exists(), reason="reference build not installed in baseline/_ref")

PORT = [48790]


def run_mix(master_impl, peer_impls, extra=()):
    PORT[0] += 1
    port = PORT[0]
    master = subprocess.Popen([sys.executable, str(SCRIPT), "master", master_impl, str(port)], stdin=subprocess.PIPE, stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True)
    try:
        assert "MASTER UP" in master.stdout.readline()
        world = len(peer_impls)
        peers = [subprocess.Popen([sys.executable, str(SCRIPT), "peer", impl, str(port), str(r), str(world), *extra], stdout=subprocess.PIPE,
                                  stderr=subprocess.STDOUT, text=True) for r, impl in enumerate(peer_impls)]
        outs = []
        for p in peers:
            try:
                out, _ = p.communicate(timeout=180)
            except subprocess.TimeoutExpired:
                for q in peers:
                    q.kill()
                raise
            outs.append((p.returncode, out))
        for rc, out in outs:
            assert rc == 0 and "PEER OK" in out, out[-3000:]
        hashes = {line for _, out in outs for line in out.splitlines() if line.startswith("QHASH")}
        assert len(hashes) == 1, f"quantised result differs between peers: {hashes}"
    finally:
        master.stdin.close()
        try:
            master.wait(timeout=20)
        except subprocess.TimeoutExpired:
            master.kill()


@pytest.mark.parametrize("master_impl,peer_impls", [
    ("ref", ["ours", "ours"]),          # our peers under the reference master
    ("ours", ["ref", "ref"]),           # reference peers under our master
    ("ours", ["ours", "ref"]),          # mixed ring
    ("ref", ["ref", "ours", "ours"]),   # mixed ring of three under the reference master
], ids=["refmaster-ours", "ourmaster-ref", "ourmaster-mixed", "refmaster-mixed3"])
def test_mixed_run(master_impl, peer_impls):
    run_mix(master_impl, peer_impls)


def test_topology_optimisation_in_a_mixed_ring():
    """The bandwidth-benchmark protocol (C2B hello, busy answer, timed blast) between the two implementations, followed by a
    reduction on the re-ordered ring."""
    run_mix("ours", ["ours", "ref"], extra=("topology",))
