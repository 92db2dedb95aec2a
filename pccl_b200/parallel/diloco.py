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
            for attempt in range(64):  # every failed attempt has re-established the ring; something is wrong if that never helps
                try:
                    self.last_info = self.comm.ext_diloco_outer_step(self.params, inner, self.buf, lr=self.lr, mu=self.mu,
                                                                     first=self.steps == 0, tag=tag)
                    self.steps += 1
                    return
                except PCCLError as e:
                    if e.result == Result.TOO_FEW_PEERS:
                        break  # alone: apply the local pseudo-gradient below
                    if e.result != Result.RANK_CONNECTION_LOST or attempt == 63:
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
    """One-step-delayed ("streaming") DiLoCo: the all-reduce of outer step t runs while the inner phase of step t+1
    computes; its result is applied at the end of step t+1.

    Follows the reference's protocol (python/examples/nanogpt_diloco/async_diloco.py):
      * membership (:419-431): `admit_peers()` asks `are_peers_pending()`; only when somebody waits is the reduction in
        flight awaited and `update_topology()` called, followed by an ordinary shared-state synchronisation;
      * pipeline insertion of a late joiner (:551-580): in the step after a topology change the peers that hold a
        reduction result apply it and THEN re-send the shared state (SEND_ONLY) while the joiner, which has no result to
        apply, receives it (RECEIVE_ONLY) -- so the joiner starts its next inner phase from the state that includes the
        update it could not take part in.
    Different from the reference: on CUDA tensors nothing goes through the host and no thread is needed -- the reduction
    is enqueued stream-ordered on a side stream (`all_reduce(..., enqueue=True)`) and overlaps with the inner phase's
    kernels; the verdict is taken when the next outer step needs the result."""

    def __init__(self, comm: Communicator, params: torch.Tensor, lr: float = 0.7, momentum: float = 0.9,
                 quantization: Optional[QuantizationOptions] = None):
        assert params.dtype == torch.float32 and params.is_contiguous()
        self.comm = comm
        self.params = params
        self.lr, self.mu = lr, momentum
        self.quantization = quantization
        self.buf = torch.zeros_like(params)
        self.grads = [torch.empty_like(params), torch.empty_like(params)]  # double buffer
        self.in_flight = None  # (kind, handle or None, buffer index)
        self.side = torch.cuda.Stream(device=params.device) if params.is_cuda else None
        self.steps = 0
        self.applied = 0
        self.topology_updated = True  # a peer that has just connected went through a topology round

    # ---- pieces -------------------------------------------------------------------------------------------------
    def _apply(self, grad: torch.Tensor) -> None:
        if self.params.is_cuda:
            outer_nesterov_(self.params, self.buf, grad, None, self.lr, self.mu, self.applied == 0)
        else:
            if self.applied == 0:
                self.buf.copy_(grad)
            else:
                self.buf.mul_(self.mu).add_(grad)
            self.params.sub_(self.lr * (grad + self.mu * self.buf))
        self.applied += 1

    def _launch(self, idx: int) -> None:
        grad = self.grads[idx]
        try:
            if grad.is_cuda:
                self.side.wait_stream(torch.cuda.current_stream(grad.device))  # the pseudo-gradient was computed there
                self.comm.all_reduce(grad, grad, op=ReduceOp.AVG, tag=idx, quantization_options=self.quantization, stream=self.side,
                                     enqueue=True)
                self.in_flight = ("stream", None, idx)
            else:
                h = self.comm.all_reduce_async(grad, grad, op=ReduceOp.AVG, tag=idx, quantization_options=self.quantization)
                self.in_flight = ("async", h, idx)
        except PCCLError as e:
            if e.result == Result.TOO_FEW_PEERS:
                self.in_flight = ("local", None, idx)  # alone: the local pseudo-gradient is the average
            elif e.result == Result.RANK_CONNECTION_LOST:
                self.in_flight = None  # a peer vanished right at the launch: this update is skipped
            else:
                raise

    def _await(self) -> Optional[torch.Tensor]:
        """Waits for the reduction in flight.  Returns the averaged pseudo-gradient, or None when there was none or it
        was lost to a peer failure (DiLoCo tolerates a skipped outer update)."""
        if self.in_flight is None:
            return None
        kind, handle, idx = self.in_flight
        self.in_flight = None
        if kind == "stream":
            try:
                self.comm.ext_wait_stream_ops(idx)
            except PCCLError as e:
                if e.result != Result.RANK_CONNECTION_LOST:
                    raise
                return None
            torch.cuda.current_stream(self.params.device).wait_stream(self.side)
        elif kind == "async":
            if not handle.wait()[0]:
                return None
        return self.grads[idx]

    # ---- the protocol ---------------------------------------------------------------------------------------------
    def admit_peers(self, sync_shared_state=None) -> bool:
        """Top of an outer iteration.  When peers wait to join: finish the reduction in flight (its result is applied by the
        coming outer_step), accept them, and run the ordinary shared-state synchronisation `sync_shared_state(strategy)`.
        Returns whether the topology changed."""
        from .._pccl import Attribute, SharedStateSyncStrategy
        changed = self.topology_updated
        if self.steps > 0:
            changed = False
            if self.comm.get_attribute(Attribute.GLOBAL_WORLD_SIZE) > 1 and self.comm.are_peers_pending():
                self._held = self._await_keep()
                self.comm.update_topology()
                changed = True
            elif self.comm.get_attribute(Attribute.GLOBAL_WORLD_SIZE) < 2:
                self._held = self._await_keep()
                self.comm.update_topology()  # alone: let anybody in who knocks
                changed = self.comm.get_attribute(Attribute.GLOBAL_WORLD_SIZE) > 1
        if changed and sync_shared_state is not None and self.comm.get_attribute(Attribute.PEER_GROUP_WORLD_SIZE) > 1:
            sync_shared_state(SharedStateSyncStrategy.ENFORCE_POPULAR)
        self.topology_updated = changed
        return changed

    _held = None  # (result or None, True) of a reduction that had to be awaited before a topology change

    def _await_keep(self):
        had = self.in_flight is not None
        return (self._await(), had)

    def outer_step(self, inner: torch.Tensor, sync_shared_state=None, run_step: Optional[int] = None) -> bool:
        """End of an outer iteration (after the inner phase).  Applies the PREVIOUS step's averaged pseudo-gradient, inserts
        a late joiner when the topology changed (second synchronisation, SEND_ONLY / RECEIVE_ONLY), resets the inner
        parameters to the outer ones and launches this step's reduction.  Returns False when the previous reduction was
        lost to a peer failure (that update is skipped).
        run_step: the RUN's outer-step counter (part of the shared state, so a late joiner has it too); it picks the tag
        and the buffer of this step's reduction, which must be the same on every peer.  Default: this peer's own count."""
        from .._pccl import Attribute, SharedStateSyncStrategy
        idx = (self.steps if run_step is None else run_step) % 2
        grad = self.grads[idx]
        # this step's pseudo-gradient, against the outer state as it was during the inner phase
        if self.params.is_cuda:
            pseudo_gradient(grad, self.params, inner)
        else:
            torch.sub(self.params, inner, out=grad)
        if self._held is not None:
            result, had_previous = self._held
            self._held = None
        else:
            had_previous = self.in_flight is not None
            result = self._await()
        ok = result is not None or not had_previous
        multi = self.comm.get_attribute(Attribute.PEER_GROUP_WORLD_SIZE) > 1
        if self.steps > 0:  # a member of the running pipeline
            if result is not None:
                self._apply(result)
            if self.topology_updated and multi and sync_shared_state is not None:
                sync_shared_state(SharedStateSyncStrategy.SEND_ONLY)  # the joiner eavesdrops on the update it missed
        elif self.topology_updated and multi and sync_shared_state is not None and self._joined_running_group:
            sync_shared_state(SharedStateSyncStrategy.RECEIVE_ONLY)  # we are the joiner: take the state after that update
        inner.copy_(self.params)
        self._launch(idx)
        self.steps += 1
        return ok

    _joined_running_group = False

    def mark_joined_running_group(self, yes: bool = True) -> None:
        """Tell a freshly connected peer that the group it joined is past its first outer step (its first shared-state
        synchronisation received data): it then takes part in the insertion protocol as the receiving side."""
        self._joined_running_group = yes

    def flush(self) -> None:
        """Waits for the reduction in flight and applies it (before a checkpoint or at the end of the run)."""
        result = self._await()
        if result is not None:
            self._apply(result)
