"""Collective backends over flat arenas.

``TorchCollectives`` is the portable backend (gloo on CPU -- what the tests in this
container run -- and NCCL on GPUs; it is also the "NCCL-only baseline" the fused path is
measured against).  ``parallel.peer.PeerCollectives`` is the product: hand-written
sm_100a kernels that read and write peer GPUs' arenas directly over NVLink.

Every method works on a *view of an arena obtained from ``alloc``* so that the peer
backend can translate the view into an offset valid on every rank.
"""
from __future__ import annotations

from typing import Dict, Optional

import torch
import torch.distributed as dist

from ..ops import reference as ref


class Hyper:
    """Optimizer hyper-parameters passed to the fused reduce+update kernels."""

    def __init__(self, kind: str = "adam", lr: float = 1e-3, beta1: float = 0.9, beta2: float = 0.999,
                 eps: float = 1e-8, weight_decay: float = 0.0, momentum: float = 0.9):
        self.kind, self.lr, self.beta1, self.beta2 = kind, lr, beta1, beta2
        self.eps, self.weight_decay, self.momentum = eps, weight_decay, momentum


class TorchCollectives:
    name = "torch"

    def __init__(self, world):
        self.world = world
        self.size = world.size
        self.device = world.device

    # -- memory ---------------------------------------------------------------------
    def alloc(self, numel: int, dtype: torch.dtype) -> torch.Tensor:
        return torch.zeros(numel, dtype=dtype, device=self.device)

    # -- collectives ----------------------------------------------------------------
    def all_reduce_avg_(self, view: torch.Tensor, algo: str = "auto") -> None:
        if self.size > 1:
            dist.all_reduce(view)
            view.mul_(1.0 / self.size)

    def all_reduce_update_(self, grads: torch.Tensor, params: torch.Tensor, state: Dict[str, torch.Tensor],
                           shadow: Optional[torch.Tensor], hyper: Hyper, step: int, algo: str = "auto",
                           wire: Optional[torch.Tensor] = None) -> None:
        """Average ``grads`` over ranks and apply one optimizer step to ``params`` (flat views).
        Portable formulation: all-reduce, then the flat update as separate passes.  With ``wire`` (a bf16 staging
        view) the exchange carries bf16 values: each rank's contribution is rounded once, the sum is fp32."""
        if wire is not None:
            wire.copy_(grads)
            grads = wire.float()
        self.all_reduce_avg_(grads, algo)
        if hyper.kind == "adam":
            ref.adam_step(params, grads, state["exp_avg"], state["exp_avg_sq"], step, hyper.lr, hyper.beta1,
                          hyper.beta2, hyper.eps, hyper.weight_decay)
        else:
            ref.sgd_step(params, grads, state["momentum_buffer"], step, hyper.lr, hyper.momentum,
                         hyper.weight_decay)
        if shadow is not None:
            shadow.copy_(params)

    def broadcast_(self, view: torch.Tensor, root: int = 0) -> None:
        if self.size > 1:
            dist.broadcast(view, src=root)

    def all_reduce_sum_i64(self, value: torch.Tensor) -> torch.Tensor:
        if self.size > 1:
            dist.all_reduce(value)
        return value

    def all_gather_slices_(self, view: torch.Tensor) -> None:
        """Each rank owns slice ``rank`` of ``view`` (equal split); make the full view valid everywhere."""
        if self.size > 1:
            parts = list(view.chunk(self.size))
            dist.all_gather(parts, parts[self.world.rank].clone())

    def owns_full_state(self, algo: str) -> bool:
        return True

    def barrier(self) -> None:
        if self.size > 1:
            dist.barrier()

    def sync_stream(self) -> None:
        pass

    def close(self) -> None:
        pass


def make_collectives(world, algo: str = "auto"):
    """Pick the backend: peer-memory kernels on multi-GPU CUDA worlds unless ``algo == 'nccl'``."""
    if world.device.type == "cuda" and algo != "nccl":
        from .peer import PeerCollectives
        return PeerCollectives(world)
    return TorchCollectives(world)
