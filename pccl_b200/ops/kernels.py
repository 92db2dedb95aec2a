"""Single-kernel operations on torch tensors.

Every function here launches hand-written sm_100a kernels from libpccl.so through the C ABI in
include/pccl_kernels.h.  There is no PyTorch fallback: on a CUDA tensor a missing library or a
failed launch raises.  Host tensors go to the C++ host kernels (the oracles).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from .. import _lib
from .._lib import lib
from .._pccl import DataType, QuantizationAlgorithm, ReduceOp


def _check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed with CUDA error {rc}")


def _stream_ptr(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _hash(t: torch.Tensor, hash_type: int) -> int:
    assert t.is_contiguous()
    out = C.c_uint64()
    nbytes = t.numel() * t.element_size()
    if t.is_cuda:
        with torch.cuda.device(t.device):
            torch.cuda.current_stream().synchronize()
            _check(lib().pcclKernelHashSync(t.data_ptr(), nbytes, hash_type, C.byref(out)), "pcclKernelHashSync")
    else:
        _check(lib().pcclExtHashBuffer(t.data_ptr(), nbytes, 0, hash_type, C.byref(out)), "pcclExtHashBuffer")
    return out.value


def simplehash(t: torch.Tensor) -> int:
    """Shared-state content hash (value-compatible with the reference's simplehash)."""
    return _hash(t, 0)


def crc32c(t: torch.Tensor) -> int:
    """CRC32C (Castagnoli) of the tensor's bytes."""
    return _hash(t, 1)


def reduce_(dst: torch.Tensor, src: torch.Tensor, op: ReduceOp = ReduceOp.SUM) -> torch.Tensor:
    """dst (op)= src, element-wise, in place."""
    assert dst.is_contiguous() and src.is_contiguous() and dst.dtype == src.dtype and dst.numel() == src.numel()
    dt = DataType.from_torch_dtype(dst.dtype).value
    if dst.is_cuda:
        with torch.cuda.device(dst.device):
            _check(lib().pcclKernelReduce(dst.data_ptr(), src.data_ptr(), dst.numel(), dt, op.value, _stream_ptr(dst)), "pcclKernelReduce")
    else:
        fn = lib().pcclHostReduce
        fn.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int]
        _check(fn(dst.data_ptr(), src.data_ptr(), dst.numel(), dt, op.value), "pcclHostReduce")
    return dst


def scale_inv_(dst: torch.Tensor, divisor: int) -> torch.Tensor:
    """dst /= divisor in the element type (AVG finalisation)."""
    dt = DataType.from_torch_dtype(dst.dtype).value
    if dst.is_cuda:
        with torch.cuda.device(dst.device):
            _check(lib().pcclKernelScaleInv(dst.data_ptr(), dst.numel(), dt, divisor, _stream_ptr(dst)), "pcclKernelScaleInv")
    else:
        fn = lib().pcclHostScaleInv
        fn.argtypes = [C.c_void_p, C.c_size_t, C.c_int, C.c_uint64]
        _check(fn(dst.data_ptr(), dst.numel(), dt, divisor), "pcclHostScaleInv")
    return dst


def _code_dtype(algo: QuantizationAlgorithm) -> int:
    return DataType.FLOAT8_E4M3.value if algo == QuantizationAlgorithm.BLOCK_FP8 else DataType.UINT8.value


def block_quantize(x: torch.Tensor, algo: QuantizationAlgorithm) -> torch.Tensor:
    """fp32 -> [per-128-element records][1 byte per element].  Returns a uint8 tensor."""
    assert x.dtype == torch.float32 and x.is_contiguous()
    assert algo in (QuantizationAlgorithm.BLOCK_MIN_MAX, QuantizationAlgorithm.BLOCK_FP8)
    n = x.numel()
    nbytes = lib().pcclKernelQuantizedBytes(n, algo.value)
    out = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
    if x.is_cuda:
        with torch.cuda.device(x.device):
            _check(lib().pcclKernelQuantize(x.data_ptr(), out.data_ptr(), n, algo.value, _stream_ptr(x)), "pcclKernelQuantize")
    else:
        fn = lib().pcclHostQuantize
        fn.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p]
        _check(fn(x.data_ptr(), out.data_ptr(), n, algo.value, _code_dtype(algo), None), "pcclHostQuantize")
    return out


def block_dequantize_reduce(dst: torch.Tensor, q: torch.Tensor, algo: QuantizationAlgorithm, op: Optional[ReduceOp] = None) -> torch.Tensor:
    """dst (op)= dequant(q); op None stores."""
    assert dst.dtype == torch.float32 and dst.is_contiguous() and q.dtype == torch.uint8
    opv = -1 if op is None else op.value
    if dst.is_cuda:
        with torch.cuda.device(dst.device):
            _check(lib().pcclKernelDequantizeReduce(dst.data_ptr(), q.data_ptr(), dst.numel(), algo.value, opv, _stream_ptr(dst)),
                   "pcclKernelDequantizeReduce")
    else:
        fn = lib().pcclHostDequantizeReduce
        fn.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_int, C.c_void_p, C.c_int]
        _check(fn(dst.data_ptr(), q.data_ptr(), dst.numel(), algo.value, _code_dtype(algo), None, opv), "pcclHostDequantizeReduce")
    return dst


class LocalArena:
    """A one-rank arena: control page + staging area in ordinary device memory.  Lets the
    all-reduce kernels run with world size 1 (single GPU tests, the one-peer DiLoCo step)."""

    def __init__(self, device: torch.device, staging_bytes: int):
        self.control = int(lib().pcclKernelP2PControlBytes())
        self.capacity = (staging_bytes + 4095) // 4096 * 4096 + 4096
        self.buf = torch.zeros(self.control + self.capacity, dtype=torch.uint8, device=device)
        self.words = torch.zeros(16, dtype=torch.int32).pin_memory()  # [0] abort flag, [1] status
        self.epoch = 0


def local_all_reduce(src: torch.Tensor, dst: torch.Tensor, arena: LocalArena, op: ReduceOp = ReduceOp.SUM,
                     quant: QuantizationAlgorithm = QuantizationAlgorithm.NONE, algorithm: int = 0, max_blocks: int = 0) -> int:
    """Runs the fused all-reduce kernel with world size 1 on the current stream.  Returns the
    algorithm that ran (1 one-shot, 2 two-shot).  With one rank SUM/AVG/MIN/MAX return the input
    (through the quantiser when one is selected)."""
    assert src.is_cuda and dst.is_cuda and src.is_contiguous() and dst.is_contiguous()
    args = _lib.KernelP2PArgs()
    args.arena[0] = arena.buf.data_ptr()
    args.rank, args.world = 0, 1
    args.flag_slot = 0
    arena.epoch += 1
    args.flag_base = arena.epoch * 16
    args.data_offset = arena.control
    args.data_capacity = arena.capacity
    args.src, args.dst = src.data_ptr(), dst.data_ptr()
    args.count = src.numel()
    args.dtype = DataType.from_torch_dtype(src.dtype).value
    args.op = op.value
    args.quant_algorithm = quant.value
    args.algorithm = algorithm
    args.max_blocks = max_blocks
    args.src_in_arena = 0
    args.abort_flag = arena.words.data_ptr()
    args.status = arena.words.data_ptr() + 4
    args.timeout_cycles = 4_000_000_000  # ~2 s: a lost peer or a kernel bug fails fast
    ran = C.c_int()
    with torch.cuda.device(src.device):
        _check(lib().pcclKernelAllReduceP2P(C.byref(args), C.byref(ran), _stream_ptr(src)), "pcclKernelAllReduceP2P")
    return ran.value


def local_outer_step(params: torch.Tensor, inner: torch.Tensor, momentum: torch.Tensor, arena: LocalArena, lr: float, mu: float,
                     first: bool, max_blocks: int = 0) -> None:
    """The fused DiLoCo outer-step kernel (pseudo-gradient -> fp32 AVG all-reduce -> Nesterov update, ONE launch)
    with world size 1 on the current stream.  With one rank the average is the local pseudo-gradient."""
    for t in (params, inner, momentum):
        assert t.is_cuda and t.is_contiguous() and t.dtype == torch.float32 and t.numel() == params.numel()
    args = _lib.KernelP2PArgs()
    args.arena[0] = arena.buf.data_ptr()
    args.rank, args.world = 0, 1
    args.flag_slot = 0
    arena.epoch += 1
    args.flag_base = arena.epoch * 16
    args.data_offset = arena.control
    args.data_capacity = arena.capacity
    args.count = params.numel()
    args.max_blocks = max_blocks
    args.abort_flag = arena.words.data_ptr()
    args.status = arena.words.data_ptr() + 4
    args.timeout_cycles = 4_000_000_000
    step = _lib.KernelOuterStepC()
    step.inner_in = step.inner_out = inner.data_ptr()
    step.params, step.momentum = params.data_ptr(), momentum.data_ptr()
    step.lr, step.mu, step.first = lr, mu, 1 if first else 0
    with torch.cuda.device(params.device):
        _check(lib().pcclKernelOuterStepP2P(C.byref(args), C.byref(step), _stream_ptr(params)), "pcclKernelOuterStepP2P")


class EmulatedWorld:
    """`world` ranks on ONE GPU, for tests: every rank has its own arena (control page + staging area) in this GPU's
    memory, and a collective runs as ONE cooperative kernel whose grid is (blocks, world) -- kernels of ranks that wait
    for each other must not be separate launches on one GPU.  The kernels are the production ones (same bodies, same
    barriers and flags); only the launch differs.  Keep `max_blocks * world` below the SM count."""

    def __init__(self, device: torch.device, world: int, staging_bytes: int):
        assert 1 <= world <= _lib.K_MAX_RANKS
        self.world = world
        self.device = device
        self.control = int(lib().pcclKernelP2PControlBytes())
        self.capacity = (staging_bytes + 4095) // 4096 * 4096 + 4096
        self.arenas = [torch.zeros(self.control + 2 * self.capacity, dtype=torch.uint8, device=device) for _ in range(world)]
        self.words = [torch.zeros(64, dtype=torch.int32).pin_memory() for _ in range(world)]
        self.epoch = 0

    def symmetric_view(self, rank: int, numel: int, dtype: torch.dtype, second: bool = False) -> torch.Tensor:
        """A tensor that lives in the rank's arena (the emulation of symmetric memory): first or second staging half."""
        off = self.control + (self.capacity if second else 0)
        nbytes = numel * torch.empty(0, dtype=dtype).element_size()
        assert nbytes <= self.capacity
        return self.arenas[rank][off:off + nbytes].view(dtype)

    def _args(self, rank: int, count: int, dtype: torch.dtype, flag_slot: int):
        a = _lib.KernelP2PArgs()
        for i in range(self.world):
            a.arena[i] = self.arenas[i].data_ptr()
        a.rank, a.world = rank, self.world
        a.flag_slot = flag_slot
        a.flag_base = self.epoch * 1024
        a.data_offset = self.control
        a.data_capacity = self.capacity
        a.count = count
        a.abort_flag = self.words[rank].data_ptr()
        a.status = self.words[rank].data_ptr() + 4
        a.timeout_cycles = 6_000_000_000  # ~3 s: a kernel bug fails instead of hanging the box
        return a

    def status(self):
        return [int(w[1]) for w in self.words]

    def all_reduce(self, srcs, dsts, op: ReduceOp = ReduceOp.SUM, quant: QuantizationAlgorithm = QuantizationAlgorithm.NONE,
                   algorithm: int = 0, max_blocks: int = 8, symmetric: str = "", no_trailing_barrier: bool = False,
                   want_hash: bool = False, flag_slot: int = 0):
        """symmetric: "" (ordinary buffers, staged), "in_place" (srcs[r] is symmetric_view(r) and is the output),
        "out_of_place" (srcs[r] first half, dsts[r] second half of the arena).  Returns (algorithm, raw hash sums or None)."""
        assert len(srcs) == len(dsts) == self.world and max_blocks * self.world <= 128
        self.epoch += 1
        for w in self.words:
            w.zero_()
        arr = (_lib.KernelP2PArgs * self.world)()
        hashes = torch.zeros(self.world, 2, dtype=torch.int64, device=self.device) if want_hash else None
        for r in range(self.world):
            s, d = srcs[r], dsts[r]
            a = self._args(r, s.numel(), s.dtype, flag_slot)
            a.src, a.dst = s.data_ptr(), d.data_ptr()
            a.dtype = DataType.from_torch_dtype(s.dtype).value
            a.op = op.value
            a.quant_algorithm = quant.value
            a.algorithm = algorithm
            a.max_blocks = max_blocks
            a.no_trailing_barrier = 1 if no_trailing_barrier else 0
            if symmetric:
                assert s.data_ptr() == self.arenas[r].data_ptr() + self.control
                a.src_in_arena = a.dst_in_arena = 1
                a.red_offset = self.control + (self.capacity if symmetric == "out_of_place" else 0)
                assert d.data_ptr() == self.arenas[r].data_ptr() + a.red_offset
            if want_hash:
                a.hash_out = hashes[r].data_ptr()
            arr[r] = a
        ran = C.c_int()
        with torch.cuda.device(self.device):
            _check(lib().pcclKernelEmulatedAllReduceP2P(arr, self.world, C.byref(ran), _stream_ptr(srcs[0])), "pcclKernelEmulatedAllReduceP2P")
        return ran.value, hashes

    def outer_step(self, params, inners, momenta, lr: float, mu: float, first: bool, max_blocks: int = 8, want_hash: bool = False):
        assert max_blocks * self.world <= 128
        self.epoch += 1
        for w in self.words:
            w.zero_()
        arr = (_lib.KernelP2PArgs * self.world)()
        steps = (_lib.KernelOuterStepC * self.world)()
        hashes = torch.zeros(self.world, 2, dtype=torch.int64, device=self.device) if want_hash else None
        for r in range(self.world):
            a = self._args(r, params[r].numel(), torch.float32, 0)
            a.max_blocks = max_blocks
            if want_hash:
                a.hash_out = hashes[r].data_ptr()
            arr[r] = a
            st = _lib.KernelOuterStepC()
            st.inner_in = st.inner_out = inners[r].data_ptr()
            st.params, st.momentum = params[r].data_ptr(), momenta[r].data_ptr()
            st.lr, st.mu, st.first = lr, mu, 1 if first else 0
            steps[r] = st
        with torch.cuda.device(self.device):
            _check(lib().pcclKernelEmulatedOuterStepP2P(arr, steps, self.world, _stream_ptr(params[0])), "pcclKernelEmulatedOuterStepP2P")
        return hashes


def mset64(t: torch.Tensor) -> int:
    """Order-independent content hash (csrc/kernels/mset_formula.hpp), the hash type the collective kernels can produce
    as a by-product."""
    return _hash(t, 2)


def mset64_finalize(raw: int, nbytes: int) -> int:
    fn = lib().pcclExtMsetFinalize
    fn.restype = C.c_uint64
    fn.argtypes = [C.c_uint64, C.c_uint64]
    return int(fn(raw & 0xFFFFFFFFFFFFFFFF, nbytes))
