"""Data-parallel gradient averaging over a Communicator (reference: python/examples/nanogptddp/train_pccl.py:472-534:
flatten gradients -> all-reduce -> scatter back, retry on failure, drop the step when alone)."""
from __future__ import annotations

from typing import Iterable, Optional

import torch

from .._pccl import Communicator, PCCLError, QuantizationOptions, ReduceOp, Result


class GradientAverager:
    def __init__(self, comm: Communicator, params: Iterable[torch.nn.Parameter], quantization: Optional[QuantizationOptions] = None,
                 symmetric: bool = False):
        self.comm = comm
        self.params = [p for p in params if p.requires_grad]
        total = sum(p.numel() for p in self.params)
        dev = self.params[0].device
        if symmetric and dev.type == "cuda":
            self.flat = comm.ext_symmetric_empty(total, torch.float32)  # reduced in place without staging
        else:
            self.flat = torch.empty(total, dtype=torch.float32, device=dev)
        self.quantization = quantization

    def average(self, tag: int = 0, max_retries: int = 16) -> bool:
        """Averages `.grad` of all parameters over the peer group.  Returns False when this peer is alone
        (the caller may skip the optimiser step, as the reference example does)."""
        off = 0
        for p in self.params:
            n = p.numel()
            self.flat[off:off + n].copy_(p.grad.reshape(-1) if p.grad is not None else torch.zeros(n, device=self.flat.device))
            off += n
        for _ in range(max_retries):
            try:
                self.comm.all_reduce(self.flat, self.flat, op=ReduceOp.AVG, tag=tag, quantization_options=self.quantization)
                break
            except PCCLError as e:
                if e.result == Result.TOO_FEW_PEERS:
                    return False
                if e.result != Result.RANK_CONNECTION_LOST:
                    raise
        else:
            return False
        off = 0
        for p in self.params:
            n = p.numel()
            if p.grad is not None:
                p.grad.copy_(self.flat[off:off + n].view_as(p.grad))
            off += n
        return True
