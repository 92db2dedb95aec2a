"""DiLoCo outer step on flat fp32 parameter buffers.

The reference ships DiLoCo as example scripts that move everything through the CPU
(python/examples/nanogpt_diloco/sync_diloco.py:390-540, sync_diloco_concat.py:311,501-503: pseudo-gradient =
outer - inner on the host, `all_reduce_multiple_with_retry(AVG)`, torch SGD(lr 0.7, nesterov), copy
back).  Here the same step runs where the parameters live: one kernel for the pseudo-gradient, the
fused NVLink all-reduce (or the TCP ring for remote peers), one kernel for the Nesterov update
that also refreshes the inner copy, and the content hash used by the next shared-state sync.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from .. import ops
from .._lib import lib
from .._pccl import (Communicator, PCCLError, QuantizationOptions, ReduceOp, Result)


def _rc(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed with CUDA error {rc}")


def pseudo_gradient(out: torch.Tensor, outer: torch.Tensor, inner: torch.Tensor) -> torch.Tensor:
    assert out.is_cuda and out.dtype == torch.float32 and out.is_contiguous()
    fn = lib().pcclKernelPseudoGrad
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
    with torch.cuda.device(out.device):
        _rc(fn(out.data_ptr(), outer.data_ptr(), inner.data_ptr(), out.numel(), torch.cuda.current_stream().cuda_stream), "pcclKernelPseudoGrad")
    return out


def outer_nesterov_(params: torch.Tensor, momentum: torch.Tensor, grad: torch.Tensor, inner: Optional[torch.Tensor], lr: float,
                    mu: float, first: bool) -> None:
    fn = lib().pcclKernelOuterNesterov
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t, C.c_float, C.c_float, C.c_int, C.c_void_p]
    with torch.cuda.device(params.device):
        _rc(fn(params.data_ptr(), momentum.data_ptr(), grad.data_ptr(), inner.data_ptr() if inner is not None else None, params.numel(), lr,
               mu, 1 if first else 0, torch.cuda.current_stream().cuda_stream), "pcclKernelOuterNesterov")


class LocalOuterStep:
    """Outer step of a single peer (world size 1) through the same kernels as the multi-peer path with one rank:
    `step` is the fused kernel (pseudo-gradient, reduction, Nesterov update in ONE launch) plus the content hash,
    `step_three_pass` the separate-kernels form it replaces (kept for comparison and as numerical cross-check)."""

    def __init__(self, params: torch.Tensor, lr: float = 0.7, momentum: float = 0.9):
        assert params.is_cuda and params.dtype == torch.float32 and params.is_contiguous()
        self.params = params
        self.lr, self.mu = lr, momentum
        self.buf = torch.zeros_like(params)
        self.grad = None  # scratch of the three-pass form, allocated on first use
        self.reduced = None
        self.arena = ops.LocalArena(params.device, params.numel() * 4)
        self.steps = 0
        self.launches_per_step = 2  # fused outer step, hash

    def step(self, inner: torch.Tensor, want_hash: bool = True) -> int:
        ops.local_outer_step(self.params, inner, self.buf, self.arena, self.lr, self.mu, self.steps == 0)
        self.steps += 1
        return ops.simplehash(self.params) if want_hash else 0

    def step_three_pass(self, inner: torch.Tensor, want_hash: bool = True) -> int:
        if self.grad is None:
            self.grad = torch.empty_like(self.params)
            self.reduced = torch.empty_like(self.params)
        pseudo_gradient(self.grad, self.params, inner)
        ops.local_all_reduce(self.grad, self.reduced, self.arena, op=ReduceOp.AVG)
        outer_nesterov_(self.params, self.buf, self.reduced, inner, self.lr, self.mu, self.steps == 0)
        self.steps += 1
        return ops.simplehash(self.params) if want_hash else 0


class DiLoCo:
    """Synchronous DiLoCo over a Communicator.

    `outer_step(inner)` averages the pseudo-gradients of all peers of the group and applies the
    outer optimiser.  Peers that drop mid-step are handled by the retry loop; a peer that ends
    up alone applies its own pseudo-gradient (what the reference examples do).
    """

    def __init__(self, comm: Communicator, params: torch.Tensor, lr: float = 0.7, momentum: float = 0.9,
                 quantization: Optional[QuantizationOptions] = None):
        assert params.dtype == torch.float32 and params.is_contiguous()
        self.comm = comm
        self.params = params
        self.lr, self.mu = lr, momentum
        self.quantization = quantization
        self.buf = torch.zeros_like(params)
        self.grad = None  # scratch for the quantised / single-peer path, allocated on first use
        self.steps = 0
        self.last_info = None

    def outer_step(self, inner: torch.Tensor, tag: int = 0) -> None:
        """One collective call (`Communicator.ext_diloco_outer_step`): between GPUs of one NVLink domain the whole step
        is a single kernel per peer.  With wire quantisation the three stages run separately (the quantised
        reduction needs the pseudo-gradient in memory)."""
        if self.quantization is None:
            while True:
                try:
                    self.last_info = self.comm.ext_diloco_outer_step(self.params, inner, self.buf, lr=self.lr, mu=self.mu,
                                                                     first=self.steps == 0, tag=tag)
                    self.steps += 1
                    return
                except PCCLError as e:
                    if e.result == Result.TOO_FEW_PEERS:
                        break  # alone: apply the local pseudo-gradient below
                    if e.result != Result.RANK_CONNECTION_LOST:
                        raise
        if self.grad is None:
            self.grad = torch.empty_like(self.params)
        if self.params.is_cuda:
            pseudo_gradient(self.grad, self.params, inner)
        else:
            torch.sub(self.params, inner, out=self.grad)
        while self.quantization is not None:
            try:
                self.last_info = self.comm.all_reduce(self.grad, self.grad, op=ReduceOp.AVG, tag=tag,
                                                      quantization_options=self.quantization)
                break
            except PCCLError as e:
                if e.result == Result.TOO_FEW_PEERS:
                    break  # alone: keep the local pseudo-gradient
                if e.result != Result.RANK_CONNECTION_LOST:
                    raise
        if self.params.is_cuda:
            outer_nesterov_(self.params, self.buf, self.grad, inner, self.lr, self.mu, self.steps == 0)
        else:
            if self.steps == 0:
                self.buf.copy_(self.grad)
            else:
                self.buf.mul_(self.mu).add_(self.grad)
            self.params.sub_(self.lr * (self.grad + self.mu * self.buf))
            inner.copy_(self.params)
        self.steps += 1


class AsyncDiLoCo:
    """One-step-delayed ("streaming") DiLoCo: the all-reduce of outer step t runs while the inner
    phase of step t+1 computes; its result is applied at the end of step t+1.
    Reference: python/examples/nanogpt_diloco/async_diloco.py:405-640 (reduction on a background thread,
    `are_peers_pending` gate before `update_topology`)."""

    def __init__(self, comm: Communicator, params: torch.Tensor, lr: float = 0.7, momentum: float = 0.9,
                 quantization: Optional[QuantizationOptions] = None):
        assert params.dtype == torch.float32 and params.is_contiguous()
        self.comm = comm
        self.params = params
        self.lr, self.mu = lr, momentum
        self.quantization = quantization
        self.buf = torch.zeros_like(params)
        self.grads = [torch.empty_like(params), torch.empty_like(params)]  # double buffer
        self.in_flight = None  # (handle, buffer index)
        self.steps = 0
        self.applied = 0

    def _apply(self, grad: torch.Tensor, inner: Optional[torch.Tensor]) -> None:
        if self.params.is_cuda:
            outer_nesterov_(self.params, self.buf, grad, inner, self.lr, self.mu, self.applied == 0)
        else:
            if self.applied == 0:
                self.buf.copy_(grad)
            else:
                self.buf.mul_(self.mu).add_(grad)
            self.params.sub_(self.lr * (grad + self.mu * self.buf))
            if inner is not None:
                inner.copy_(self.params)
        self.applied += 1

    def outer_step(self, inner: torch.Tensor) -> bool:
        """Launches the reduction of this step's pseudo-gradient and applies the previous one.
        Returns False when the previous reduction was lost to a peer failure (it is skipped)."""
        idx = self.steps % 2
        grad = self.grads[idx]
        if self.params.is_cuda:
            pseudo_gradient(grad, self.params, inner)
        else:
            torch.sub(self.params, inner, out=grad)
        ok = True
        previous = self.in_flight
        if previous is not None:
            handle, pidx = previous
            success, status, _ = handle.wait()
            if success:
                self._apply(self.grads[pidx], None)
            else:
                ok = False  # dropped: DiLoCo tolerates a skipped outer update
        try:
            self.in_flight = (self.comm.all_reduce_async(grad, grad, op=ReduceOp.AVG, tag=idx, quantization_options=self.quantization), idx)
        except PCCLError as e:
            if e.result != Result.TOO_FEW_PEERS:
                raise
            self.in_flight = None
            self._apply(grad, None)  # alone: apply the local pseudo-gradient right away
        inner.copy_(self.params)
        self.steps += 1
        return ok

    def flush(self) -> None:
        """Waits for the reduction in flight and applies it (call before checkpointing / topology changes)."""
        if self.in_flight is not None:
            handle, pidx = self.in_flight
            if handle.wait()[0]:
                self._apply(self.grads[pidx], None)
            self.in_flight = None
