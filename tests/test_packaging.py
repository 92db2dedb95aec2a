"""`pip wheel .` / `pip install`: the wheel carries the native library, the master executable and the console script, and
the installed copy works on its own, away from the source tree (reference: python/framework/setup.py, which runs CMake and
drops libpccl into the package)."""
import os
import subprocess
import sys
import zipfile
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent

USE_IT = r"""
import sys, threading
import numpy as np
import pccl_b200 as pccl
assert pccl.__file__.startswith(sys.argv[1]), pccl.__file__
with pccl.MasterNode("127.0.0.1:0") as master:
    master.run()
    out = {}
    def peer(rank):
        comm = pccl.Communicator(f"127.0.0.1:{master.port}")
        comm.connect(n_attempts=20)
        while comm.get_attribute(pccl.Attribute.GLOBAL_WORLD_SIZE) < 2:
            comm.update_topology()
        x = np.full(1000, rank + 1.0, dtype=np.float32)
        comm.all_reduce(x, x, op=pccl.ReduceOp.SUM)
        out[rank] = float(x[0])
        comm.destroy()
    ts = [threading.Thread(target=peer, args=(r,)) for r in range(2)]
    [t.start() for t in ts]; [t.join() for t in ts]
assert out == {0: 3.0, 1: 3.0}, out
print("installed copy works")
"""


def test_wheel_builds_offline_and_the_installed_copy_works(tmp_path):
    wheels, site = tmp_path / "wheels", tmp_path / "site"
    pip = [sys.executable, "-m", "pip", "--disable-pip-version-check"]
    proc = subprocess.run(pip + ["wheel", "--no-index", "--no-build-isolation", "--no-deps", "-w", str(wheels), str(REPO)],
                          stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1800)
    assert proc.returncode == 0, proc.stdout[-4000:]
    wheel = next(wheels.glob("pccl_b200-*.whl"))
    names = zipfile.ZipFile(wheel).namelist()
    assert "pccl_b200/lib/libpccl.so" in names and "pccl_b200/bin/ccoip_master" in names
    assert any(n.endswith("entry_points.txt") for n in names)
    assert "pccl_master" in zipfile.ZipFile(wheel).read(next(n for n in names if n.endswith("entry_points.txt"))).decode()
    proc = subprocess.run(pip + ["install", "--no-index", "--no-deps", "--target", str(site), str(wheel)], stdout=subprocess.PIPE,
                          stderr=subprocess.STDOUT, text=True, timeout=600)
    assert proc.returncode == 0, proc.stdout[-4000:]
    env = {k: v for k, v in os.environ.items() if k != "PYTHONPATH"}
    env["PYTHONPATH"] = str(site)
    proc = subprocess.run([sys.executable, "-c", USE_IT, str(site)], cwd=str(tmp_path), stdout=subprocess.PIPE, stderr=subprocess.STDOUT,
                          text=True, timeout=300, env=env)
    assert proc.returncode == 0 and "installed copy works" in proc.stdout, proc.stdout[-4000:]
    master = site / "pccl_b200" / "bin" / "ccoip_master"
    os.chmod(master, 0o755)  # pip keeps the mode bits of package data only when they were set in the wheel
    proc = subprocess.run([str(master), "--help"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=30, cwd=str(tmp_path))
    assert "usage: ccoip_master" in proc.stdout, proc.stdout
