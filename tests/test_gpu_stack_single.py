"""Whole stack on ONE GPU: two peers share the device, so the NVLink plane must refuse (kernels of
two ranks that wait for each other may not share a GPU) and the TCP ring with device staging runs."""
import pytest

from test_gpu_multi import run_peers

pytestmark = pytest.mark.gpu


def test_two_peers_one_gpu_fall_back_to_tcp():
    results = run_peers(2, "allreduce", devices=["cuda:0", "cuda:0"])
    for r in results:
        for c in r["cases"]:
            assert c["ok"], c
            assert c["plane"] == "tcp_ring", c


def test_shared_state_two_peers_one_gpu():
    for r in run_peers(2, "shared_state", devices=["cuda:0", "cuda:0"]):
        assert r["ok"], r


def test_outer_step_two_peers_one_gpu():
    """Device tensors, TCP plane: the outer step takes its two-pass form (pseudo-gradient kernel into the library's scratch,
    staged ring reduction, Nesterov kernel once the verdict is in)."""
    results = run_peers(2, "outer_step", devices=["cuda:0", "cuda:0"], env={"PCCL_ARENA_SLOT_MB": "64", "PCCL_SYMMETRIC_MB": "128"})
    for r in results:
        assert r["ok"], r
        assert all(c["plane"] == "tcp_ring" for c in r["cases"]), r
    assert len({c["hash"] for r in results for c in r["cases"] if c["id"] == "outer_step_1"}) == 1
