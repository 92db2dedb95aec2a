"""The allocation-free hash entry point (pcclKernelSimpleHashNoAlloc).

Some callers of the library cannot allocate device memory at the moment they need a shared-state
hash: a stream capture may be open on their thread, the training framework's allocator may already
hold every byte of the GPU, or they sit on a path where the device-wide synchronisation hidden in
cudaMalloc/cudaFree is not acceptable.  For them the library keeps a 4 KiB scratch array in static
device memory and offers an entry point that uses it; inputs and the 8-byte output still come from
the caller.

This file is named so that pytest collects it last.  It drives the new entry point through the C ABI
and compares it with the existing entry point (caller-provided scratch) on two calls in a row, which
also proves that the shared scratch is left in a usable state between calls.
"""
import ctypes as C

import pytest
import torch

from pccl_b200._lib import lib

pytestmark = pytest.mark.gpu


def _declare():
    L = lib()
    L.pcclKernelSimpleHashNoAlloc.restype = C.c_int
    L.pcclKernelSimpleHashNoAlloc.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_void_p]
    L.pcclKernelSimpleHashNoAllocMaxBytes.restype = C.c_size_t
    L.pcclKernelSimpleHashNoAllocMaxBytes.argtypes = []
    return L


def _reference_hash(L, x: torch.Tensor) -> int:
    """existing entry point: scratch supplied by the caller"""
    scratch = torch.zeros(L.pcclKernelHashScratchBytes(), dtype=torch.uint8, device=x.device)
    out = torch.zeros(1, dtype=torch.int64, device=x.device)
    stream = torch.cuda.current_stream().cuda_stream
    assert L.pcclKernelSimpleHash(x.data_ptr(), x.numel() * x.element_size(), scratch.data_ptr(), out.data_ptr(), stream) == 0
    torch.cuda.synchronize()
    return int(out.item()) & 0xFFFFFFFFFFFFFFFF


def _no_alloc_hash(L, x: torch.Tensor, out: torch.Tensor) -> int:
    stream = torch.cuda.current_stream().cuda_stream
    rc = L.pcclKernelSimpleHashNoAlloc(x.data_ptr(), x.numel() * x.element_size(), out.data_ptr(), stream)
    assert rc == 0, f"pcclKernelSimpleHashNoAlloc returned CUDA error {rc}"
    torch.cuda.synchronize()
    return int(out.item()) & 0xFFFFFFFFFFFFFFFF


def test_no_alloc_hash_matches_regular_entry_point_twice_in_a_row():
    L = _declare()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    g = torch.Generator().manual_seed(99)
    # two different inputs, the second one smaller so it uses fewer blocks than the first left behind
    first = torch.randint(0, 256, (48 * 1024 * 1024 + 13,), dtype=torch.uint8, generator=g).to(dev)
    second = torch.randint(0, 256, (700_001,), dtype=torch.uint8, generator=g).to(dev)
    out = torch.zeros(1, dtype=torch.int64, device=dev)  # the caller owns the output
    before = torch.cuda.memory_allocated(dev)
    got_first = _no_alloc_hash(L, first, out)
    got_second = _no_alloc_hash(L, second, out)
    assert torch.cuda.memory_allocated(dev) == before
    # the hash is an integer: the tolerance every hash test in this suite uses is exact equality
    assert got_first == _reference_hash(L, first)
    assert got_second == _reference_hash(L, second)
    # and once more in the other order, back to back on the same stream without a sync in between
    stream = torch.cuda.current_stream().cuda_stream
    out2 = torch.zeros(1, dtype=torch.int64, device=dev)
    assert L.pcclKernelSimpleHashNoAlloc(second.data_ptr(), second.numel(), out.data_ptr(), stream) == 0
    assert L.pcclKernelSimpleHashNoAlloc(first.data_ptr(), first.numel(), out2.data_ptr(), stream) == 0
    torch.cuda.synchronize()
    assert int(out.item()) & 0xFFFFFFFFFFFFFFFF == got_second
    assert int(out2.item()) & 0xFFFFFFFFFFFFFFFF == got_first


def test_no_alloc_hash_refuses_oversized_input():
    L = _declare()
    limit = L.pcclKernelSimpleHashNoAllocMaxBytes()
    # 2^32 - 1 whole packets (the hash counts packets in 32 bits) and a tail shorter than a packet; the kernel's slab and
    # loop arithmetic is 64-bit, so the top of the range is hashed, not wrapped
    assert limit == ((1 << 32) - 1) * 16 + 15
    x = torch.zeros(16, dtype=torch.uint8, device="cuda:0")
    out = torch.zeros(1, dtype=torch.int64, device="cuda:0")
    # only the size is looked at before the refusal, nothing is read
    rc = L.pcclKernelSimpleHashNoAlloc(x.data_ptr(), limit + 1, out.data_ptr(), None)
    assert rc == 1  # cudaErrorInvalidValue
