"""Loads libpccl.so (built in-tree by `make` / `__graft_entry__.build()`) and declares the C ABI.

Counterpart of the reference's cffi loader (reference: python/framework/pccl/_loader.py:12-27 and the
generated python/framework/pccl/_cdecls.py); ctypes is used here so the package needs nothing beyond
the standard library.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
import sys
from pathlib import Path

_PKG_DIR = Path(__file__).resolve().parent
_REPO_DIR = _PKG_DIR.parent
LIB_PATH = _PKG_DIR / "lib" / "libpccl.so"
MASTER_BIN = _PKG_DIR / "bin" / "ccoip_master"


def build_native(force: bool = False, jobs: int | None = None) -> None:
    """Runs `make` in the repository root.  All CUDA sources are compiled for sm_100a."""
    if LIB_PATH.exists() and MASTER_BIN.exists() and not force:
        return
    jobs = jobs or max(2, os.cpu_count() or 2)
    cmd = ["make", "-C", str(_REPO_DIR), f"-j{jobs}", "lib", "master"]
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if proc.returncode != 0:
        sys.stderr.write(proc.stdout[-8000:])
        raise RuntimeError("building libpccl.so failed")


class IPv4(C.Structure):
    _fields_ = [("data", C.c_uint8 * 4)]


class IPv6(C.Structure):
    _fields_ = [("data", C.c_uint8 * 16)]


class InetAddress(C.Structure):
    _fields_ = [("protocol", C.c_int), ("ipv4", IPv4), ("ipv6", IPv6)]


class SocketAddress(C.Structure):
    _fields_ = [("inet", InetAddress), ("port", C.c_uint16)]


class CommCreateParams(C.Structure):
    _fields_ = [
        ("master_address", SocketAddress),
        ("peer_group", C.c_uint32),
        ("p2p_connection_pool_size", C.c_uint32),
        ("use_explicit_p2p_addresses", C.c_bool),
        ("advertised_p2p_address", SocketAddress),
        ("advertised_shared_state_address", SocketAddress),
        ("advertised_benchmark_address", SocketAddress),
        ("internal_p2p_listen_port", C.c_uint16),
        ("internal_shared_state_listen_port", C.c_uint16),
        ("internal_benchmark_listen_port", C.c_uint16),
    ]


class ReduceInfoC(C.Structure):
    _fields_ = [("local_world_size", C.c_uint32), ("tx_bytes", C.c_uint64), ("rx_bytes", C.c_uint64)]


class OperandDescriptorC(C.Structure):
    _fields_ = [("datatype", C.c_int), ("distribution_hint", C.c_int)]


class QuantizationOptionsC(C.Structure):
    _fields_ = [("quantized_datatype", C.c_int), ("algorithm", C.c_int)]


class ReduceDescriptorC(C.Structure):
    _fields_ = [
        ("count", C.c_size_t),
        ("op", C.c_int),
        ("tag", C.c_uint64),
        ("src_descriptor", OperandDescriptorC),
        ("quantization_options", QuantizationOptionsC),
    ]


class ReduceOpDescriptorC(C.Structure):
    _fields_ = [("sendbuf", C.c_void_p), ("recvbuf", C.c_void_p), ("descriptor", ReduceDescriptorC)]


class AsyncReduceOpC(C.Structure):
    _fields_ = [("comm", C.c_void_p), ("tag", C.c_uint64)]


class TensorInfoC(C.Structure):
    _fields_ = [
        ("name", C.c_char_p),
        ("data", C.c_void_p),
        ("count", C.c_size_t),
        ("datatype", C.c_int),
        ("device_type", C.c_int),
        ("allow_content_inequality", C.c_bool),
    ]


class SharedStateC(C.Structure):
    _fields_ = [("revision", C.c_uint64), ("count", C.c_size_t), ("infos", C.POINTER(TensorInfoC))]


class SharedStateSyncInfoC(C.Structure):
    _fields_ = [("tx_bytes", C.c_uint64), ("rx_bytes", C.c_uint64)]


class BuildInfoC(C.Structure):
    _fields_ = [("has_cuda_support", C.c_bool)]


class ExtReduceStatsC(C.Structure):
    _fields_ = [("plane", C.c_int), ("device_ms", C.c_float), ("seq_nr", C.c_uint64), ("kernel_launches", C.c_uint32),
                ("leased", C.c_uint32), ("phase_ns", C.c_uint64 * 8),
                ("has_content_hash", C.c_uint32), ("content_hash", C.c_uint64 * 2)]


class ExtOuterStepDescriptorC(C.Structure):
    _fields_ = [("params", C.c_void_p), ("inner", C.c_void_p), ("momentum", C.c_void_p), ("count", C.c_uint64),
                ("lr", C.c_float), ("mu", C.c_float), ("first_step", C.c_int), ("tag", C.c_uint64)]


class KernelOuterStepC(C.Structure):
    _fields_ = [("inner_in", C.c_void_p), ("params", C.c_void_p), ("momentum", C.c_void_p), ("inner_out", C.c_void_p),
                ("lr", C.c_float), ("mu", C.c_float), ("first", C.c_int)]


K_MAX_RANKS = 16


class KernelP2PArgs(C.Structure):
    _fields_ = [
        ("arena", C.c_void_p * K_MAX_RANKS),
        ("rank", C.c_uint32),
        ("world", C.c_uint32),
        ("flag_slot", C.c_uint32),
        ("flag_base", C.c_uint64),
        ("data_offset", C.c_uint64),
        ("data_capacity", C.c_uint64),
        ("src", C.c_void_p),
        ("dst", C.c_void_p),
        ("count", C.c_uint64),
        ("dtype", C.c_int),
        ("op", C.c_int),
        ("quant_algorithm", C.c_int),
        ("algorithm", C.c_int),
        ("max_blocks", C.c_int),
        ("src_in_arena", C.c_int),
        ("abort_flag", C.c_void_p),
        ("status", C.c_void_p),
        ("timeout_cycles", C.c_uint64),
        ("red_offset", C.c_uint64),
        ("dst_in_arena", C.c_int),
        ("no_trailing_barrier", C.c_int),
        ("stamps", C.c_void_p),
        ("hash_out", C.c_void_p),
    ]


_lib = None


def lib() -> C.CDLL:
    """The loaded library with argument / result types declared.  Loading is loud: a missing
    shared object is an error, never a silent fallback."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        if os.environ.get("PCCL_B200_NO_AUTOBUILD", "0") == "1":
            raise OSError(f"{LIB_PATH} is missing; run `make` in {_REPO_DIR}")
        build_native()
    L = C.CDLL(str(LIB_PATH), mode=C.RTLD_GLOBAL)
    vp, u64, sz, i32 = C.c_void_p, C.c_uint64, C.c_size_t, C.c_int

    def sig(name, restype, *argtypes):
        fn = getattr(L, name)
        fn.restype = restype
        fn.argtypes = list(argtypes)

    sig("pcclInit", i32)
    sig("pcclCreateCommunicator", i32, C.POINTER(CommCreateParams), C.POINTER(vp))
    sig("pcclGetAttribute", i32, vp, i32, C.POINTER(i32))
    sig("pcclDestroyCommunicator", i32, vp)
    sig("pcclConnect", i32, vp)
    sig("pcclUpdateTopology", i32, vp)
    sig("pcclArePeersPending", i32, vp, C.POINTER(C.c_bool))
    sig("pcclOptimizeTopology", i32, vp)
    sig("pcclAllReduce", i32, vp, vp, C.POINTER(ReduceDescriptorC), vp, C.POINTER(ReduceInfoC))
    sig("pcclAllReduceAsync", i32, vp, vp, C.POINTER(ReduceDescriptorC), vp, C.POINTER(AsyncReduceOpC))
    sig("pcclAllReduceMultipleWithRetry", i32, C.POINTER(ReduceOpDescriptorC), sz, vp, C.POINTER(ReduceInfoC), i32)
    sig("pcclAwaitAsyncReduce", i32, C.POINTER(AsyncReduceOpC), C.POINTER(ReduceInfoC))
    sig("pcclSynchronizeSharedState", i32, vp, C.POINTER(SharedStateC), i32, C.POINTER(SharedStateSyncInfoC))
    sig("pcclCreateMaster", i32, SocketAddress, C.POINTER(vp))
    sig("pcclRunMaster", i32, vp)
    sig("pcclInterruptMaster", i32, vp)
    sig("pcclMasterAwaitTermination", i32, vp)
    sig("pcclDestroyMaster", i32, vp)
    sig("pcclGetBuildInfo", i32, C.POINTER(BuildInfoC))
    sig("pcclExtAllReduceOnStream", i32, vp, vp, C.POINTER(ReduceDescriptorC), vp, vp, i32, C.POINTER(ReduceInfoC))
    sig("pcclExtTrustContentHashes", i32, vp, i32)
    sig("pcclExtHashKernelsSkipped", i32, vp, C.POINTER(u64))
    sig("pcclExtWaitStreamOps", i32, vp, u64, i32, C.POINTER(ReduceInfoC))
    sig("pcclExtGetLastReduceStats", i32, vp, u64, C.POINTER(ExtReduceStatsC))
    sig("pcclExtMasterGetPort", i32, vp, C.POINTER(C.c_uint16))
    sig("pcclExtHashBuffer", i32, vp, sz, i32, i32, C.POINTER(u64))
    sig("pcclExtSymmetricAlloc", i32, vp, sz, C.POINTER(vp))
    sig("pcclExtSymmetricFree", i32, vp, vp)
    sig("pcclExtDataTypeSize", sz, i32)
    sig("pcclExtDiLoCoOuterStep", i32, C.POINTER(ExtOuterStepDescriptorC), vp, C.POINTER(ReduceInfoC))
    # kernels
    sig("pcclKernelHashScratchBytes", sz)
    sig("pcclKernelSimpleHash", i32, vp, sz, vp, vp, vp)
    sig("pcclKernelCrc32c", i32, vp, sz, vp, vp, vp)
    sig("pcclKernelHashSync", i32, vp, sz, i32, C.POINTER(u64))
    sig("pcclKernelReduce", i32, vp, vp, sz, i32, i32, vp)
    sig("pcclKernelScaleInv", i32, vp, sz, i32, u64, vp)
    sig("pcclKernelQuantizedBytes", sz, sz, i32)
    sig("pcclKernelQuantize", i32, vp, vp, sz, i32, vp)
    sig("pcclKernelDequantizeReduce", i32, vp, vp, sz, i32, i32, vp)
    sig("pcclKernelAllReduceP2P", i32, C.POINTER(KernelP2PArgs), C.POINTER(i32), vp)
    sig("pcclKernelP2PStagingBytes", sz, u64, i32, i32)
    sig("pcclKernelOuterStepP2P", i32, C.POINTER(KernelP2PArgs), C.POINTER(KernelOuterStepC), vp)
    sig("pcclKernelPseudoGrad", i32, vp, vp, vp, sz, vp)
    sig("pcclKernelOuterNesterov", i32, vp, vp, vp, vp, sz, C.c_float, C.c_float, i32, vp)
    sig("pcclKernelP2PControlBytes", sz)
    L.pcclInit()
    _lib = L
    return L
