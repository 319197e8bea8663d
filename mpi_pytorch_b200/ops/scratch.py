"""Per-step pool of pre-zeroed fp32 scratch for the kernels that accumulate with atomics.

The conv epilogue's BatchNorm statistics and the BatchNorm backward sums each need 2*C zeroed floats.  Zeroing
them one by one costs ~70 fill launches and ~36 memsets per ResNet-34 step (~0.25 ms of a 15.7 ms step).  The
training engine instead *arms* the pool once per step with one buffer it has just zeroed with a single kernel
(inside the captured CUDA graph, so replays re-zero it); ``take`` then hands out disjoint slices.  Outside an
armed step -- plain use of the ops, evaluation, a second pass that outgrows the pool -- ``take`` falls back to
``torch.zeros``, so correctness never depends on the pool.

No counterpart in the reference (its BatchNorm is the ATen CPU kernel, models.py:33..87 build the layers).
"""
from __future__ import annotations

from typing import Optional

import torch

_ALIGN = 64                 # floats; keeps every slice 256-byte aligned

_buf: Optional[torch.Tensor] = None
_off = 0
stats = {"pooled": 0, "fallback": 0}


def new_buffer(device, floats: int = 1 << 18) -> torch.Tensor:
    return torch.zeros(floats, dtype=torch.float32, device=device)


def arm(buf: torch.Tensor) -> None:
    """Zero ``buf`` (one launch on the current stream) and serve ``take`` calls from it until ``disarm``."""
    global _buf, _off
    buf.zero_()
    _buf, _off = buf, 0


def disarm() -> None:
    global _buf, _off
    _buf, _off = None, 0


def armed() -> bool:
    return _buf is not None


def take(n: int, device) -> tuple:
    """-> (zeroed fp32 tensor of n elements, came_from_pool)."""
    global _off
    if _buf is not None and _buf.device == torch.device(device) and _off + n <= _buf.numel():
        t = _buf[_off:_off + n]
        _off += (n + _ALIGN - 1) // _ALIGN * _ALIGN
        stats["pooled"] += 1
        return t, True
    stats["fallback"] += 1
    return torch.zeros(n, dtype=torch.float32, device=device), False
