"""Runs the C++ test executables (tests/cpp/test_*.cpp) one test case per pytest item."""
import subprocess
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parent.parent
CPP = REPO / "tests" / "cpp"


def _cases():
    import re
    out = []
    for src in sorted(CPP.glob("test_*.cpp")):
        if "gpu" in src.stem:
            continue
        for name in re.findall(r"^TEST\((\w+)\)", src.read_text(), flags=re.M):
            out.append((src.stem, name))
    return out


@pytest.mark.parametrize("binary,case", _cases(), ids=lambda v: v)
def test_cpp(binary, case):
    exe = REPO / "build" / "tests" / binary
    assert exe.exists(), f"{exe} was not built"
    proc = subprocess.run([str(exe), case], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert proc.returncode == 0, proc.stdout[-4000:]


def test_headers_are_c99():
    exe = REPO / "build" / "tests" / "c99_compat"
    assert exe.exists(), f"{exe} was not built"
    proc = subprocess.run([str(exe)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=60)
    assert proc.returncode == 0 and "c99 ok" in proc.stdout, proc.stdout
