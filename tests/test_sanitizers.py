"""C++ tests rebuilt with AddressSanitizer+UBSan+LSan and with ThreadSanitizer (opt-in: PCCL_SANITIZE_TESTS=1).

Each flavour takes about a minute to build (`make SANITIZE=<kind> NO_CUDA=1 tests` -> build-<kind>/) and a few
minutes to run, so the default CPU run skips it; the last recorded outcome is in profiles/sanitizers.md.
Reference counterpart: cmake/testing.cmake:7-21 (`-fsanitize=address,leak,undefined`; the reference has no TSan run).
"""
import os
import subprocess
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parent.parent
BINARIES = ["test_units", "test_transport", "test_e2e_allreduce", "test_e2e_shared_state", "test_e2e_membership"]

pytestmark = pytest.mark.skipif(os.environ.get("PCCL_SANITIZE_TESTS") != "1", reason="set PCCL_SANITIZE_TESTS=1")


@pytest.fixture(scope="module", params=["address", "thread"])
def flavour(request):
    kind = request.param
    proc = subprocess.run(["make", f"SANITIZE={kind}", "NO_CUDA=1", "-j8", "tests"], cwd=REPO, stdout=subprocess.PIPE,
                          stderr=subprocess.STDOUT, text=True, timeout=1800)
    assert proc.returncode == 0, proc.stdout[-4000:]
    return kind


@pytest.mark.parametrize("binary", BINARIES)
def test_clean_under_sanitizer(flavour, binary):
    if binary == "test_units":
        pytest.skip("its fork()-based death tests print sanitizer reports on purpose; covered by the plain build")
    env = dict(os.environ, ASAN_OPTIONS="detect_leaks=1", TSAN_OPTIONS="halt_on_error=0 exitcode=66")
    proc = subprocess.run([str(REPO / f"build-{flavour}" / "tests" / binary)], stdout=subprocess.PIPE,
                          stderr=subprocess.STDOUT, text=True, timeout=1800, env=env)
    assert proc.returncode == 0, proc.stdout[-6000:]
    assert "ThreadSanitizer" not in proc.stdout and "AddressSanitizer" not in proc.stdout and "runtime error" not in proc.stdout
