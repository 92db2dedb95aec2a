"""The C, C++ and CUDA programs under examples/ are built against include/ and the in-tree library and run as real peers
(counterparts of the reference's docs/md hello-world and concurrent all-reduce tutorials)."""
import os
import re
import shutil
import subprocess
import time
from pathlib import Path

import pytest

REPO = Path(__file__).resolve().parent.parent
LIB = REPO / "pccl_b200" / "lib"
LINK = [f"-I{REPO / 'include'}", f"-L{LIB}", "-lpccl"]


def build(cmd, tmp_path, name):
    exe = tmp_path / name
    proc = subprocess.run(cmd + ["-o", str(exe)], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=600)
    assert proc.returncode == 0, proc.stdout[-4000:]
    return exe


def start_master(tmp_path):
    """The C++ example doubles as the master; port 0 lets the kernel choose and the program prints the choice."""
    exe = build(["/usr/bin/g++", "-std=c++17", "-O1", "-pthread", str(REPO / "examples" / "cpp" / "concurrent_allreduce.cpp"), *LINK, f"-Wl,-rpath,{LIB}"],
                tmp_path, "concurrent_allreduce")
    master = subprocess.Popen([str(exe), "--master", "127.0.0.1:0"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    line = master.stdout.readline()
    port = int(re.search(r"port (\d+)", line).group(1))
    return exe, master, port


def stop(master):
    master.send_signal(2)  # SIGINT: the example's handler interrupts the master, the program leaves cleanly
    assert master.wait(timeout=30) == 0


def test_c_peer_program_with_a_late_joiner(tmp_path):
    peer = build(["gcc", "-std=c99", "-pedantic", "-Wall", "-Wextra", "-Werror", str(REPO / "examples" / "c" / "peer.c"), *LINK, f"-Wl,-rpath,{LIB}"],
                 tmp_path, "peer")
    _, master, port = start_master(tmp_path)
    env = dict(os.environ, PCCL_EXAMPLE_MASTER_PORT=str(port))
    try:
        args = [str(peer), "127.0.0.1", "60", "2", "50"]  # 60 steps of at least 50 ms: three seconds for the newcomer to arrive
        procs = [subprocess.Popen(args, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env) for _ in range(2)]
        time.sleep(0.6)
        procs.append(subprocess.Popen(args, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env))
        outs = [p.communicate(timeout=120)[0] for p in procs]
        assert all(p.returncode == 0 for p in procs), outs
        finals = {re.search(r"finished: (.*)", o).group(1) for o in outs}
        assert len(finals) == 1, finals  # the late joiner ends with the same weights as the others
        got = re.search(r"received (\d+) bytes of state, continuing at step (\d+)", outs[2])
        assert got and int(got.group(1)) == 4096 * 4 + 8 and int(got.group(2)) > 0, outs[2][:500]
        assert "with 3 peers" in outs[0]
    finally:
        stop(master)


def test_cpp_concurrent_allreduce_program(tmp_path):
    exe, master, port = start_master(tmp_path)
    try:
        procs = [subprocess.Popen([str(exe), f"127.0.0.1:{port}"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for _ in range(2)]
        outs = [p.communicate(timeout=120)[0] for p in procs]
        assert all(p.returncode == 0 for p in procs), outs
        for o in outs:
            plain = re.search(r"plain fp32: 8 operations .* (\d+) bytes sent, largest error ([\d.]+)", o)
            quant = re.search(r"uint8 min-max: 8 operations .* (\d+) bytes sent, largest error ([\d.]+)", o)
            assert plain and quant and "list of 8 averages done with 2 peers" in o, o
            assert float(plain.group(2)) == 0.0 and 0 < float(quant.group(2)) < 0.05
            assert int(quant.group(1)) * 3 < int(plain.group(1))  # a quarter of the bytes plus the ranges
    finally:
        stop(master)


@pytest.mark.skipif(shutil.which("nvcc") is None, reason="no CUDA toolkit on this machine")
def test_cuda_example_compiles_for_sm_100a(tmp_path):
    """Cross-compiles on the CPU box; it runs in tests/test_gpu_multi.py where there are two GPUs."""
    build_cuda_example(tmp_path)


def build_cuda_example(tmp_path):
    return build(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", str(REPO / "examples" / "cuda" / "stream_allreduce.cu"), *LINK,
                  "-Xlinker", f"-rpath,{LIB}"], tmp_path, "stream_allreduce")
