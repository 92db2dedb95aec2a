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


def test_two_level_ring_with_device_tensors():
    """Four peers on this GPU pretend to be two hosts of two (PCCL_HOST_ID): device tensors are staged and reduced over the
    local ring, the ring between hosts and the local ring again."""
    hosts = [{"PCCL_HOST_ID": h} for h in ("north", "south", "north", "south")]
    small = {"PCCL_ARENA_SLOT_MB": "64", "PCCL_SYMMETRIC_MB": "128"}
    results = run_peers(4, "allreduce", devices=["cuda:0"] * 4, env=small, rank_env=hosts)
    by_case = {}
    for r in results:
        for c in r["cases"]:
            assert c["ok"], c
            by_case.setdefault(c["id"], set()).add(c["hash"])
    assert {c["plane"] for r in results for c in r["cases"]} == {"tcp_hierarchical"}
    assert all(len(h) == 1 for h in by_case.values()), "peers disagree bit-wise"
    stepped = run_peers(4, "outer_step", devices=["cuda:0"] * 4, env=small, rank_env=hosts)
    assert all(r["ok"] for r in stepped), stepped
    assert {c["plane"] for r in stepped for c in r["cases"]} == {"tcp_hierarchical"}
    assert len({c["hash"] for r in stepped for c in r["cases"] if c["id"] == "outer_step_1"}) == 1
