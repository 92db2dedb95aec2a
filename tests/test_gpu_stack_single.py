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
