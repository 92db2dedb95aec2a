"""Multi-rank arithmetic of the peer-to-peer kernels on ONE GPU.

The production kernel bodies run as one cooperative kernel whose grid is (blocks, ranks); every rank has its own arena
in this GPU's memory (`ops.EmulatedWorld`).  So a single-GPU box checks real reductions -- W inputs folded into one
result, cross-rank barriers, chunk ownership, symmetric and in-place operands -- against a plain PyTorch fold in a wider
type, not just `out == x` at world size 1.  (The reference has no device kernels to compare with; its host kernels are
ccoip/src/cpp/reduce_kernels.cpp:56-215 and the oracle here restates their arithmetic.)"""
import pytest
import torch

import pccl_b200 as pccl
from pccl_b200 import ops

pytestmark = pytest.mark.gpu

DTYPES = [torch.uint8, torch.int8, torch.int16, torch.uint16, torch.int32, torch.uint32, torch.int64, torch.uint64,
          torch.float16, torch.bfloat16, torch.float32, torch.float64]
OPS = [pccl.ReduceOp.SUM, pccl.ReduceOp.AVG, pccl.ReduceOp.PROD, pccl.ReduceOp.MAX, pccl.ReduceOp.MIN]


def dev():
    if not torch.cuda.is_available():
        pytest.skip("needs a GPU")
    return torch.device("cuda:0")


def make(rank, n, dtype, seed):
    g = torch.Generator().manual_seed(seed * 100 + rank)
    if dtype.is_floating_point:
        return (torch.rand(n, generator=g) * 1.5 + 0.25).to(dtype)
    return torch.randint(1, 3, (n,), generator=g).to(torch.int64).to(dtype)


def fold(xs, op, dtype):
    wide = torch.float64 if dtype.is_floating_point else torch.int64
    acc = xs[0].to(wide)
    for x in xs[1:]:
        y = x.to(wide)
        if op in (pccl.ReduceOp.SUM, pccl.ReduceOp.AVG):
            acc = acc + y
        elif op == pccl.ReduceOp.PROD:
            acc = acc * y
        elif op == pccl.ReduceOp.MAX:
            acc = torch.maximum(acc, y)
        else:
            acc = torch.minimum(acc, y)
    if op == pccl.ReduceOp.AVG:
        acc = acc / len(xs) if dtype.is_floating_point else torch.div(acc, len(xs), rounding_mode="trunc")
    return acc


def close(got, want, dtype):
    if not dtype.is_floating_point:
        return torch.equal(got.cpu().to(torch.int64), want)
    rtol = {torch.float32: 2e-6, torch.float64: 1e-13, torch.bfloat16: 2e-2, torch.float16: 3e-3}[dtype]
    return torch.allclose(got.cpu().double(), want, rtol=rtol, atol=rtol)


@pytest.mark.parametrize("world", [2, 3, 4])
def test_every_type_and_operator(world):
    d = dev()
    n = 4099  # not a multiple of anything: partial packets, chunks of unequal size
    w = ops.EmulatedWorld(d, world, 1 << 20)
    for algorithm in (1, 2):
        for dtype in DTYPES:
            for op in OPS:
                xs = [make(r, n, dtype, 7) for r in range(world)]
                srcs = [x.to(d) for x in xs]
                dsts = [torch.zeros_like(s) for s in srcs]
                ran, _ = w.all_reduce(srcs, dsts, op=op, algorithm=algorithm, max_blocks=4)
                assert ran == algorithm and w.status() == [0] * world
                want = fold(xs, op, dtype)
                for r in range(world):
                    assert close(dsts[r], want, dtype), (algorithm, dtype, op, r)
                # every rank holds the same bits
                for r in range(1, world):
                    assert torch.equal(dsts[r].view(torch.uint8), dsts[0].view(torch.uint8))


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("n", [3, 1000, 70_001, 1_000_003])
def test_sizes_and_in_place_on_ordinary_buffers(world, n):
    d = dev()
    w = ops.EmulatedWorld(d, world, n * 4 + 4096)
    xs = [make(r, n, torch.float32, 11) for r in range(world)]
    want = fold(xs, pccl.ReduceOp.AVG, torch.float32)
    for algorithm in (0, 2):
        bufs = [x.to(d) for x in xs]
        w.all_reduce(bufs, bufs, op=pccl.ReduceOp.AVG, algorithm=algorithm, max_blocks=16)
        assert w.status() == [0] * world
        for b in bufs:
            assert close(b, want, torch.float32)


@pytest.mark.parametrize("world", [2, 3, 4])
@pytest.mark.parametrize("mode", ["in_place", "out_of_place"])
def test_symmetric_operands(world, mode):
    """Operands that live in the arenas themselves.  Small in-place ones used to race in the one-shot kernel (a rank
    overwrote its packet before a slower peer had read it); sizes around the one-shot limit and counts that are not a
    multiple of four are covered on purpose."""
    d = dev()
    for n in (5, 4099, 65_536, 65_541, 300_001):
        w = ops.EmulatedWorld(d, world, n * 4 + 64)
        xs = [make(r, n, torch.float32, 13) for r in range(world)]
        want = fold(xs, pccl.ReduceOp.SUM, torch.float32)
        srcs = [w.symmetric_view(r, n, torch.float32) for r in range(world)]
        dsts = srcs if mode == "in_place" else [w.symmetric_view(r, n, torch.float32, second=True) for r in range(world)]
        for rep in range(3):  # repeated: a race shows up as an occasional double count
            for r in range(world):
                srcs[r].copy_(xs[r])
            ran, _ = w.all_reduce(srcs, dsts, op=pccl.ReduceOp.SUM, max_blocks=8, symmetric=mode)
            assert w.status() == [0] * world
            for r in range(world):
                assert close(dsts[r], want, torch.float32), (n, rep, r, ran)
            if mode == "out_of_place":
                for r in range(world):
                    assert torch.equal(srcs[r].cpu(), xs[r]), "the input of an out-of-place reduction must survive"


@pytest.mark.parametrize("world", [2, 4])
@pytest.mark.parametrize("algo", [pccl.QuantizationAlgorithm.BLOCK_MIN_MAX, pccl.QuantizationAlgorithm.BLOCK_FP8])
def test_quantised_two_shot(world, algo):
    d = dev()
    n = 200_003
    w = ops.EmulatedWorld(d, world, n * 2 + 65536)
    xs = [torch.randn(n, generator=torch.Generator().manual_seed(50 + r)) for r in range(world)]
    want = fold(xs, pccl.ReduceOp.AVG, torch.float32)
    for in_place in (False, True):
        srcs = [x.to(d) for x in xs]
        dsts = srcs if in_place else [torch.zeros_like(s) for s in srcs]
        w.all_reduce(srcs, dsts, op=pccl.ReduceOp.AVG, quant=algo, max_blocks=8)
        assert w.status() == [0] * world
        tol = (0.03 if algo == pccl.QuantizationAlgorithm.BLOCK_MIN_MAX else 0.12) * want.abs().max().item()
        for r in range(world):
            assert (dsts[r].cpu().double() - want).abs().max().item() <= tol
            assert torch.equal(dsts[r], dsts[0]), "owners adopt what the others decode: identical bits everywhere"


@pytest.mark.parametrize("world", [2, 3])
def test_fused_content_hash_of_the_result(world):
    """The two-shot kernel hashes the chunk it reduces and the ranks swap partial sums: every rank ends up with the mset64
    digest of the COMPLETE result, equal to a stand-alone hash pass over the output and to the host twin."""
    d = dev()
    for n, dtype in ((1_000_003, torch.float32), (70_001, torch.bfloat16), (17, torch.float32)):
        w = ops.EmulatedWorld(d, world, n * 4 + 4096)
        srcs = [make(r, n, dtype, 17).to(d) for r in range(world)]
        dsts = [torch.zeros_like(s) for s in srcs]
        _, raw = w.all_reduce(srcs, dsts, op=pccl.ReduceOp.SUM, algorithm=2, max_blocks=8, want_hash=True)
        nbytes = n * dsts[0].element_size()
        for r in range(world):
            fused = ops.mset64_finalize(int(raw[r, 0].item()), nbytes)
            assert fused == ops.mset64(dsts[r]) == ops.mset64(dsts[r].cpu()), (n, dtype, r)


@pytest.mark.parametrize("world", [2, 4])
def test_outer_step_as_one_kernel(world):
    d = dev()
    n = 300_007
    w = ops.EmulatedWorld(d, world, n * 4 + 4096)
    start = torch.randn(n, generator=torch.Generator().manual_seed(1))
    params = [start.clone().to(d) for _ in range(world)]
    moms = [torch.full((n,), 7.0, device=d) for _ in range(world)]
    ref_p, ref_m = start.clone().double(), torch.zeros(n, dtype=torch.float64)
    lr, mu = 0.7, 0.9
    for k in range(2):
        drifts = [0.01 * torch.randn(n, generator=torch.Generator().manual_seed(100 + 10 * k + r)) for r in range(world)]
        inners = [(ref_p.float() - dr).to(d) for dr in drifts]
        for r in range(world):
            params[r].copy_(ref_p.float())
        g = sum((ref_p.float() - (ref_p.float() - dr)).double() for dr in drifts) / world
        ref_m = g.clone() if k == 0 else mu * ref_m + g
        ref_p = ref_p.float().double() - lr * (g + mu * ref_m)
        raw = w.outer_step(params, inners, moms, lr, mu, first=k == 0, max_blocks=8, want_hash=True)
        assert w.status() == [0] * world
        for r in range(world):
            assert torch.allclose(params[r].cpu().double(), ref_p, rtol=1e-5, atol=1e-6)
            assert torch.allclose(moms[r].cpu().double(), ref_m, rtol=1e-5, atol=1e-6)
            assert torch.equal(inners[r], params[r])
            assert torch.equal(params[r], params[0])
            assert ops.mset64_finalize(int(raw[r, 0].item()), n * 4) == ops.mset64(params[r])
            assert ops.mset64_finalize(int(raw[r, 1].item()), n * 4) == ops.mset64(moms[r])
        ref_p = params[0].cpu().double()
        ref_m = moms[0].cpu().double()


def test_mset64_device_equals_host():
    d = dev()
    for n in (0, 1, 15, 16, 17, 4096, 1_000_003):
        x = torch.randint(0, 255, (n,), dtype=torch.uint8)
        assert ops.mset64(x.to(d)) == ops.mset64(x), n
    # order independence is the point: a permutation of whole packets changes the value (the index is part of each term)
    x = torch.randint(0, 255, (64,), dtype=torch.uint8)
    y = torch.cat([x[16:32], x[:16], x[32:]])
    assert ops.mset64(x) != ops.mset64(y)


@pytest.mark.parametrize("world", [2, 3])
def test_zero_point_scale_on_the_device(world):
    """Zero-point/scale quantisation (uint8 codes, one record per 128 elements) in the fused kernel.  The reference's anchor
    (ccoip/tests/end_to_end/test_all_reduce.cpp:326-399): constants 42 and 43 must come out as exactly 85."""
    d = dev()
    n = 70_003
    w = ops.EmulatedWorld(d, world, n * 2 + 65536)
    consts = [42.0, 43.0, 44.0][:world]
    srcs = [torch.full((n,), c, device=d) for c in consts]
    dsts = [torch.zeros_like(s) for s in srcs]
    w.all_reduce(srcs, dsts, op=pccl.ReduceOp.SUM, quant=pccl.QuantizationAlgorithm.ZERO_POINT_SCALE, max_blocks=8)
    assert w.status() == [0] * world
    for out in dsts:
        assert torch.equal(out, torch.full_like(out, sum(consts)))
    xs = [torch.randn(n, generator=torch.Generator().manual_seed(90 + r)) * 3 + r for r in range(world)]
    want = fold(xs, pccl.ReduceOp.SUM, torch.float32)
    srcs = [x.to(d) for x in xs]
    w.all_reduce(srcs, dsts, op=pccl.ReduceOp.SUM, quant=pccl.QuantizationAlgorithm.ZERO_POINT_SCALE, max_blocks=8)
    tol = 0.03 * want.abs().max().item()
    for out in dsts:
        assert (out.cpu().double() - want).abs().max().item() <= tol
        assert torch.equal(out, dsts[0])
