"""Single-GPU numerics of the sm_100a kernels against host oracles and plain PyTorch fp32."""
import ctypes as C

import pytest
import torch

import pccl_b200 as pccl
from pccl_b200 import ops
from pccl_b200._lib import lib

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


def lcg_bytes(n: int, seed: int = 42) -> torch.Tensor:
    """Byte stream of the reference's hash tests (simplehash_cuda_test.cpp:10-15): 32-bit LCG, value % 256."""
    import numpy as np
    a, c = np.uint64(1664525), np.uint64(1013904223)
    # closed form per index would need modular exponentiation; iterate in numpy chunks instead
    out = np.empty(n, dtype=np.uint8)
    s = np.uint64(seed)
    mask = np.uint64(0xFFFFFFFF)
    for i in range(n):
        s = (a * s + c) & mask
        out[i] = s % np.uint64(256)
    return torch.from_numpy(out)


@pytest.mark.parametrize("n,golden", [(1, 344386053), (4, 3765247898), (25, 3651434421)])
def test_simplehash_golden_small(n, golden):
    host = lcg_bytes(n)
    assert ops.simplehash(host) == golden
    assert ops.simplehash(host.to(DEV)) == golden


@pytest.mark.parametrize("nbytes", [16, 48, 4096, 4096 * 3 + 7, 256 * 16 * 960, 256 * 16 * 960 + 16, 10_000_019, 64 << 20])
def test_simplehash_matches_host(nbytes):
    g = torch.Generator().manual_seed(nbytes)
    host = torch.randint(0, 256, (nbytes,), dtype=torch.uint8, generator=g)
    assert ops.simplehash(host.to(DEV)) == ops.simplehash(host)


def test_simplehash_is_repeatable_and_order_sensitive():
    x = torch.randint(0, 2**31 - 1, (1 << 20,), dtype=torch.int32, device=DEV)
    h1, h2 = ops.simplehash(x), ops.simplehash(x)
    assert h1 == h2
    y = x.clone()
    y[12345], y[12346] = x[12346], x[12345]
    assert ops.simplehash(y) != h1 or x[12345] == x[12346]


@pytest.mark.parametrize("nbytes", [0, 1, 9, 511, 512, 513, 8191, 1 << 20, (1 << 20) + 333, 50_000_003])
def test_crc32c_matches_host(nbytes):
    g = torch.Generator().manual_seed(nbytes + 1)
    host = torch.randint(0, 256, (max(nbytes, 1),), dtype=torch.uint8, generator=g)[:nbytes].contiguous()
    want = ops.crc32c(host)
    if nbytes == 9:
        assert ops.crc32c(torch.tensor(list(b"123456789"), dtype=torch.uint8)) == 0xE3069283
    got = ops.crc32c(host.to(DEV)) if nbytes else want
    assert got == want


@pytest.mark.parametrize("dtype", [torch.float32, torch.float64, torch.float16, torch.bfloat16, torch.int32, torch.int64, torch.uint8, torch.int8, torch.int16])
@pytest.mark.parametrize("op", [pccl.ReduceOp.SUM, pccl.ReduceOp.MAX, pccl.ReduceOp.MIN, pccl.ReduceOp.PROD])
def test_reduce_matches_torch(dtype, op):
    n = 100_003
    g = torch.Generator().manual_seed(5)
    if dtype.is_floating_point:
        a = torch.randn(n, generator=g).to(dtype)
        b = torch.randn(n, generator=g).to(dtype)
    else:
        a = torch.randint(0, 11, (n,), generator=g).to(dtype)
        b = torch.randint(0, 11, (n,), generator=g).to(dtype)
    ref32 = {pccl.ReduceOp.SUM: torch.add, pccl.ReduceOp.MAX: torch.maximum, pccl.ReduceOp.MIN: torch.minimum, pccl.ReduceOp.PROD: torch.mul}[op]
    if dtype.is_floating_point:
        want = ref32(a.float() if dtype != torch.float64 else a, b.float() if dtype != torch.float64 else b).to(dtype)
    else:
        want = ref32(a, b)
    got = ops.reduce_(a.to(DEV), b.to(DEV), op).cpu()
    if dtype.is_floating_point:
        torch.testing.assert_close(got.float(), want.float(), rtol=1e-6 if dtype in (torch.float32, torch.float64) else 1e-2, atol=1e-6)
    else:
        assert torch.equal(got, want)


def torch_block_minmax(x: torch.Tensor):
    """plain PyTorch fp32 reference of the block min-max quantiser"""
    n = x.numel()
    pad = (-n) % 128
    xp = torch.cat([x, x[-1:].expand(pad)]) if pad else x
    blocks = xp.view(-1, 128)
    mn, mx = blocks.min(dim=1, keepdim=True).values, blocks.max(dim=1, keepdim=True).values
    # tensor / tensor: `scalar / tensor` is evaluated as reciprocal * scalar (two roundings)
    inv = torch.where(mx > mn, torch.full_like(mx, 255.0) / (mx - mn), torch.zeros_like(mx))
    codes = torch.clamp(torch.round((blocks - mn) * inv), 0, 255)
    deq = torch.addcmul(mn, codes, (mx - mn) * (1.0 / 255.0))
    return codes.view(-1)[:n].to(torch.uint8), deq.view(-1)[:n]


@pytest.mark.parametrize("n", [128, 1000, 128 * 1024 + 5, 3_000_017])
def test_block_minmax_quantiser(n):
    g = torch.Generator().manual_seed(n)
    x = torch.randn(n, generator=g) * 3 + 0.5
    q_dev = ops.block_quantize(x.to(DEV), pccl.QuantizationAlgorithm.BLOCK_MIN_MAX).cpu()
    q_host = ops.block_quantize(x, pccl.QuantizationAlgorithm.BLOCK_MIN_MAX)
    assert torch.equal(q_dev, q_host), "device and host quantisers must agree bit for bit"
    codes_ref, deq_ref = torch_block_minmax(x)
    n_blocks = (n + 127) // 128
    assert torch.equal(q_dev[n_blocks * 8:], codes_ref)
    out = torch.zeros(n, device=DEV)
    ops.block_dequantize_reduce(out, q_dev.to(DEV), pccl.QuantizationAlgorithm.BLOCK_MIN_MAX)
    torch.testing.assert_close(out.cpu(), deq_ref, rtol=0, atol=2e-6)
    assert (out.cpu() - x).abs().max() <= (x.max() - x.min()) / 255.0


@pytest.mark.parametrize("n", [128, 777, 2_000_003])
def test_block_fp8_quantiser(n):
    g = torch.Generator().manual_seed(n)
    x = torch.randn(n, generator=g) * 10
    q_dev = ops.block_quantize(x.to(DEV), pccl.QuantizationAlgorithm.BLOCK_FP8)
    q_host = ops.block_quantize(x, pccl.QuantizationAlgorithm.BLOCK_FP8)
    assert torch.equal(q_dev.cpu(), q_host)
    out_dev = torch.zeros(n, device=DEV)
    ops.block_dequantize_reduce(out_dev, q_dev, pccl.QuantizationAlgorithm.BLOCK_FP8)
    # plain PyTorch reference: scale to the e4m3 range per block, cast, scale back
    pad = (-n) % 128
    xp = torch.cat([x, torch.zeros(pad)]) if pad else x
    blocks = xp.view(-1, 128)
    amax = blocks.abs().max(dim=1, keepdim=True).values
    scale = torch.where(amax > 0, amax / 448.0, torch.ones_like(amax))
    ref = ((blocks * (1.0 / scale)).to(torch.float8_e4m3fn).float() * scale).view(-1)[:n]
    torch.testing.assert_close(out_dev.cpu(), ref, rtol=0, atol=1e-6)
    # accumulate variant
    acc = torch.ones(n, device=DEV)
    ops.block_dequantize_reduce(acc, q_dev, pccl.QuantizationAlgorithm.BLOCK_FP8, pccl.ReduceOp.SUM)
    torch.testing.assert_close(acc.cpu(), ref + 1.0, rtol=0, atol=1e-5)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16, torch.float64, torch.int32])
@pytest.mark.parametrize("n", [1, 5, 4096, 1_000_003, 16 << 20])
@pytest.mark.parametrize("algorithm", [1, 2])
def test_fused_allreduce_world1(dtype, n, algorithm):
    """World size 1 exercises staging, the per-block barrier, the owner reduction and the write-out."""
    g = torch.Generator().manual_seed(n)
    x = (torch.randn(n, generator=g) * 4).to(dtype).to(DEV)
    arena = ops.LocalArena(torch.device(DEV), n * x.element_size())
    out = torch.empty_like(x)
    for op in (pccl.ReduceOp.SUM, pccl.ReduceOp.AVG, pccl.ReduceOp.MAX):
        out.zero_()
        ran = ops.local_all_reduce(x, out, arena, op=op, algorithm=algorithm)
        torch.cuda.synchronize()
        assert ran == algorithm
        assert int(arena.words[1]) == 0
        assert torch.equal(out, x)
    # unaligned views take the element-wise path
    if n > 8:
        xs, os_ = x[1:], out[1:]
        os_.zero_()
        ops.local_all_reduce(xs, os_, arena, algorithm=algorithm)
        torch.cuda.synchronize()
        assert torch.equal(os_, xs)


@pytest.mark.parametrize("algo", [pccl.QuantizationAlgorithm.BLOCK_MIN_MAX, pccl.QuantizationAlgorithm.BLOCK_FP8])
@pytest.mark.parametrize("n", [7, 128, 100_001, 8 << 20])
def test_fused_quantised_allreduce_world1(algo, n):
    """With one rank the fused kernel returns D(Q(D(Q(x)))): the staging quantisation followed by the
    owner's re-quantisation.  Re-quantising already decoded values is the identity for min-max
    blocks whose extremes are code words, so compare against the standalone quantiser applied twice."""
    g = torch.Generator().manual_seed(n)
    x = (torch.randn(n, generator=g) * 2).to(DEV)
    arena = ops.LocalArena(torch.device(DEV), n * 2 + 4096)
    out = torch.zeros_like(x)
    ops.local_all_reduce(x, out, arena, quant=algo)
    torch.cuda.synchronize()
    assert int(arena.words[1]) == 0
    once = torch.zeros_like(x)
    ops.block_dequantize_reduce(once, ops.block_quantize(x, algo), algo)
    twice = torch.zeros_like(x)
    ops.block_dequantize_reduce(twice, ops.block_quantize(once, algo), algo)
    torch.testing.assert_close(out, twice, rtol=0, atol=1e-6)
    tol = 0.02 if algo == pccl.QuantizationAlgorithm.BLOCK_MIN_MAX else 0.08
    assert (out - x).abs().max().item() <= tol * x.abs().max().item() * 2


# ---- fused DiLoCo outer step (one launch) against plain PyTorch fp32 -----------------------------
@pytest.mark.parametrize("n", [1, 3, 4, 1000, 4099, 1 << 20, (1 << 22) + 5])
@pytest.mark.parametrize("offset", [0, 1])
def test_fused_outer_step_matches_torch(n, offset):
    from pccl_b200.parallel.diloco import LocalOuterStep
    dev = torch.device("cuda:0")
    torch.manual_seed(n)

    def place(t):  # offset 1: the tensor starts 4 bytes off a 16-byte boundary (scalar path of the kernel)
        big = torch.empty(t.numel() + offset, device=dev)
        view = big[offset:]
        view.copy_(t)
        return view

    params = place(torch.randn(n, device=dev))
    inner = place(params - 0.01 * torch.randn(n, device=dev))
    step = LocalOuterStep(params, lr=0.7, momentum=0.9)
    ref_p, ref_m = params.clone(), torch.zeros_like(params)
    for k in range(3):
        g = ref_p - inner
        ref_m = g.clone() if k == 0 else 0.9 * ref_m + g
        ref_p = ref_p - 0.7 * (g + 0.9 * ref_m)
        step.step(inner, want_hash=False)
        torch.cuda.synchronize()
        torch.testing.assert_close(step.params, ref_p, rtol=1e-5, atol=1e-6)
        torch.testing.assert_close(step.buf, ref_m, rtol=1e-5, atol=1e-6)
        assert torch.equal(inner, step.params), "inner must receive the new parameters"
        inner.sub_(0.003 * torch.randn_like(inner))  # next inner phase drifts again


def test_fused_outer_step_equals_three_pass():
    from pccl_b200.parallel.diloco import LocalOuterStep
    dev = torch.device("cuda:0")
    torch.manual_seed(5)
    n = 3_000_001
    outer = torch.randn(n, device=dev)
    inner_a = outer - 0.02 * torch.randn(n, device=dev)
    inner_b = inner_a.clone()
    a, b = LocalOuterStep(outer.clone()), LocalOuterStep(outer.clone())
    for _ in range(2):
        a.step(inner_a, want_hash=False)
        b.step_three_pass(inner_b, want_hash=False)
        torch.cuda.synchronize()
        torch.testing.assert_close(a.params, b.params, rtol=1e-6, atol=1e-7)
        torch.testing.assert_close(a.buf, b.buf, rtol=1e-6, atol=1e-7)
        torch.testing.assert_close(inner_a, inner_b, rtol=1e-6, atol=1e-7)
        inner_a.sub_(0.01)
        inner_b.sub_(0.01)


@pytest.mark.parametrize("n", [262_144, 262_148, 1_000_000, 16_000_004, 37_000_000])
def test_staging_through_the_bulk_copy_engine(n):
    """Operands of a megabyte and more are staged with cp.async.bulk (global -> shared -> global) instead of register
    copies; sizes that end in a partial tile and both SUM and AVG go through it.  World size 1: the result is the input."""
    dev = torch.device("cuda:0")
    x = torch.randn(n, device=dev)
    arena = ops.LocalArena(dev, n * 4)
    for op in (pccl.ReduceOp.SUM, pccl.ReduceOp.AVG):
        y = torch.zeros_like(x)
        assert ops.local_all_reduce(x, y, arena, op=op) == 2
        torch.cuda.synchronize()
        assert int(arena.words[1]) == 0
        assert torch.equal(x, y)
    # in place as well (the output is written only after the commit point)
    z = x.clone()
    ops.local_all_reduce(z, z, arena, op=pccl.ReduceOp.SUM)
    assert torch.equal(x, z)
