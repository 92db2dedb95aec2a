"""`pip install .` builds the native library first (reference: python/framework/setup.py:1-186 runs CMake and drops
libpccl.so into the package; here the Makefile does the same job).

Environment, as in the reference:
  PCCL_BUILD_CUDA_SUPPORT=0   build without the CUDA plane and kernels (`make NO_CUDA=1`); default: on when nvcc is found
"""
import os
import shutil
import subprocess
from pathlib import Path

from setuptools import setup
from setuptools.command.build_py import build_py

ROOT = Path(__file__).resolve().parent


class BuildNativeThenPy(build_py):
    def run(self):
        nvcc = os.environ.get("NVCC") or shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
        want_cuda = os.environ.get("PCCL_BUILD_CUDA_SUPPORT", "1") not in ("0", "OFF", "off", "false")
        args = ["make", "-j", str(os.cpu_count() or 4), "lib", "master"]
        if not (want_cuda and Path(nvcc).exists()):
            args.append("NO_CUDA=1")
        subprocess.check_call(args, cwd=ROOT)
        super().run()


setup(cmdclass={"build_py": BuildNativeThenPy})
