"""pytest configuration: markers and the one-time native build."""
import os
import subprocess
import sys
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on a B200 box)")
    config.addinivalue_line("markers", "multigpu: needs at least two CUDA devices")
    config.addinivalue_line("markers", "slow: long-running")


@pytest.fixture(scope="session", autouse=True)
def native_build():
    """Builds libpccl.so, ccoip_master and the C++ test binaries once per session."""
    have = (REPO / "pccl_b200" / "lib" / "libpccl.so").exists() and (REPO / "pccl_b200" / "bin" / "ccoip_master").exists() \
        and any((REPO / "build" / "tests").glob("test_*"))
    if have and os.environ.get("PCCL_REBUILD", "0") != "1":
        # artefacts travel with the snapshot to the GPU box; file times may not, so `make` is
        # only run when something is missing (or on request)
        yield
        return
    jobs = max(2, os.cpu_count() or 2)
    proc = subprocess.run(["make", "-C", str(REPO), f"-j{jobs}", "all"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if proc.returncode != 0:
        sys.stderr.write(proc.stdout[-6000:])
        pytest.exit("native build failed", returncode=2)
    yield


def gpu_count() -> int:
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:
        return 0
