#!/usr/bin/env python3
"""Builds the UNMODIFIED reference (jundi69/pccl) into baseline/_ref so `bench.py --impl reference`, the interop
tests and the reference's own Python suite can run.  Called by `__graft_entry__.build()`; a no-op when the install
is already there, and never fatal (the reference arm then reports "unavailable").

Why not the plain `pip install /root/reference`: the tree has no root setup.py (the package lives in
python/framework and is built by CMake from the repository root), the source tree is read-only while the build
writes into it, and the `third_party/piquant` submodule is missing from the scrape although
ccoip/internal/piquant_utils.hpp:5 includes it unconditionally.  So: copy the tree to a temporary directory, drop the
header-only stand-in from baseline/piquant_shim into third_party/piquant (the timed configurations use
pcclQuantNone / pcclQuantMinMax and never reach it), then
`pip install --no-index --no-build-isolation --no-deps --find-links /opt/wheelhouse --target baseline/_ref python/framework`.
Reference sources are not edited.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import tempfile
from pathlib import Path

REPO = Path(__file__).resolve().parent.parent
REF_SRC = Path(os.environ.get("PCCL_REFERENCE_SRC", "/root/reference"))
TARGET = REPO / "baseline" / "_ref"


def installed() -> bool:
    return (TARGET / "pccl" / "__init__.py").exists() and any((TARGET / "pccl").glob("*pccl*.so"))


def install(verbose: bool = True) -> tuple[bool, str]:
    if installed():
        return True, "already installed"
    if not (REF_SRC / "python" / "framework" / "setup.py").exists():
        return False, f"reference sources not found under {REF_SRC}"
    tmp = Path(tempfile.mkdtemp(prefix="pccl_ref_"))
    try:
        work = tmp / "src"
        shutil.copytree(REF_SRC, work, symlinks=True, ignore=shutil.ignore_patterns(".git"))
        subprocess.run(["chmod", "-R", "u+w", str(work)], check=False)
        shim_dst = work / "third_party" / "piquant"
        shim_dst.mkdir(parents=True, exist_ok=True)
        shutil.copytree(REPO / "baseline" / "piquant_shim", shim_dst, dirs_exist_ok=True)
        env = dict(os.environ)
        # the image exports a g++ that links libstdc++ statically; the distribution compiler gives a library that can
        # share a process with torch.  CUDA support off: the reference uses CUDA only to hash shared-state tensors
        # (its all-reduce is host-memory only) and a libcuda-linked build cannot be import-checked on a GPU-less box.
        env.update({"PCCL_BUILD_CUDA_SUPPORT": "OFF", "CC": "/usr/bin/gcc", "CXX": "/usr/bin/g++",
                    "CMAKE_BUILD_PARALLEL_LEVEL": str(max(2, os.cpu_count() or 2))})
        if TARGET.exists():
            shutil.rmtree(TARGET)
        cmd = [sys.executable, "-m", "pip", "install", "--no-index", "--no-build-isolation", "--no-deps", "--find-links", "/opt/wheelhouse",
               "--target", str(TARGET), "."]
        proc = subprocess.run(cmd, cwd=work / "python" / "framework", env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if proc.returncode != 0:
            if verbose:
                sys.stderr.write(proc.stdout[-4000:])
            return False, "pip install of the reference failed: " + proc.stdout.strip().splitlines()[-1][:200]
        if not installed():
            return False, "pip reported success but baseline/_ref/pccl has no libpccl"
        return True, "installed"
    finally:
        shutil.rmtree(tmp, ignore_errors=True)


if __name__ == "__main__":
    ok, why = install()
    print(("ok: " if ok else "FAILED: ") + why)
    sys.exit(0 if ok else 1)
