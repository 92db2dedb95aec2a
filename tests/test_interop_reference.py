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


def test_python_surface_is_a_superset_of_the_reference_package():
    """Every public class, enum member, method and keyword argument of the reference's Python package exists here under
    the same name (checked in a subprocess: the two packages each load a `libpccl.so`)."""
    code = r'''
import inspect, json, sys
sys.path.insert(0, sys.argv[1] + "/baseline/_ref")
import pccl as ref
sys.path.pop(0)
sys.path.insert(0, sys.argv[1])
import pccl_b200 as ours
internal = {"to_c", "ffi", "C", "load_native_module", "ModuleDummy", "PY_PCCL_DEBUG", "version", "pccl"}
problems = []
for name in dir(ref):
    if name.startswith("_") or name in internal:
        continue
    r = getattr(ref, name)
    if inspect.ismodule(r) and name != "cuda":
        continue
    if not inspect.isclass(r) and not inspect.ismodule(r):
        continue
    if getattr(r, "__module__", "pccl").split(".")[0] not in ("pccl",) and not inspect.ismodule(r):
        continue  # re-exported typing / stdlib helpers
    if not hasattr(ours, name):
        problems.append(f"missing {name}")
        continue
    o = getattr(ours, name)
    for member in dir(r):
        if member.startswith("_") or member in internal:
            continue
        if not hasattr(o, member):
            problems.append(f"missing {name}.{member}")
            continue
        rm, om = getattr(r, member), getattr(o, member)
        if inspect.isfunction(rm) and inspect.isfunction(om):
            rp, op = inspect.signature(rm).parameters, inspect.signature(om).parameters
            for p in rp:
                if p not in op:
                    problems.append(f"{name}.{member}: parameter {p} missing")
init_r, init_o = inspect.signature(ref.Communicator.__init__).parameters, inspect.signature(ours.Communicator.__init__).parameters
problems += [f"Communicator(): parameter {p} missing" for p in init_r if p not in init_o]
print(json.dumps(problems))
'''
    proc = subprocess.run([sys.executable, "-c", code, str(REPO)], stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=120)
    assert proc.returncode == 0, proc.stderr[-3000:]
    import json
    problems = json.loads(proc.stdout.strip().splitlines()[-1])
    assert problems == [], problems


@pytest.mark.parametrize("peer_impls", [["ref", "ours", "ours"], ["ours", "ref", "ours"]], ids=["ref-distributes", "ours-distributes"])
def test_large_shared_state_between_the_implementations(peer_impls):
    """72 MB in twelve tensors: our receivers fetch contiguous runs over several connections, each with the reference's plain
    request, so a reference distributor serves them without knowing; a reference receiver fetches from ours over one."""
    run_mix("ours", peer_impls, extra=("bigstate",))
