"""Flat parameter / gradient / optimizer-state arenas.

The reference keeps every parameter, gradient and Adam moment as its own small tensor
and walks them one by one (62-364 blocking ``Allreduce`` calls per step,
/root/reference/mpi_tools.py:30-37).  On B200 the unit of work is a *bucket*: a
contiguous, 256-byte aligned range of one flat arena, so that

* one kernel launch reduces (and optimises) thousands of tensors,
* the arena can be a symmetric peer-mapped allocation that other GPUs read/write
  directly over NVLink (parallel/peer.py), and
* ``model.parameters()`` / ``p.grad`` / ``optimizer.state[p]['exp_avg']`` are *views*
  into the arenas, so ``state_dict()`` still emits the stock torch layout the
  reference's ``helpers.load_checkpoint`` understands.

Buckets are filled in *reverse registration order* (head first) because that is the
order in which autograd produces gradients.
"""
from __future__ import annotations

import dataclasses
from typing import Callable, Dict, List, Optional, Sequence

import torch

ALIGN_ELEMS = 64          # 256 B in fp32: keeps every tensor 16B-vector aligned in any dtype >= 4 B/8 elems


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


@dataclasses.dataclass
class Bucket:
    index: int
    start: int                 # element offset into the arena (multiple of ALIGN_ELEMS)
    numel: int                 # padded length (multiple of ALIGN_ELEMS * world-friendly granule)
    params: List[torch.nn.Parameter]
    pending: int = 0           # grads still outstanding in the current backward
    algo: str = "auto"


@dataclasses.dataclass
class Slot:
    offset: int
    numel: int
    bucket: int


class FlatLayout:
    """Assigns each trainable parameter a slot in a flat arena and groups slots into buckets."""

    def __init__(self, params: Sequence[torch.nn.Parameter], bucket_bytes: int, world: int = 1,
                 elem_bytes: int = 4, tail_bytes: Optional[int] = None):
        """``bucket_bytes`` caps every bucket.  The buckets that complete LAST in backward (stem and first stages)
        cannot overlap with anything -- backward is over when they fire -- so the schedule ends small: the final
        bucket holds at most ``tail_bytes`` (default min(2 MB, cap / 4)), the one before at most 4x that, and only
        then the full cap applies.  A late small bucket costs a latency-bound launch instead of a bandwidth-bound
        one (32 MB over 8 GPUs: ~120 us exposed; 2 MB: ~25 us)."""
        self.params = [p for p in params if p.requires_grad]
        self.world = max(1, world)
        granule = ALIGN_ELEMS * self.world          # every bucket splits evenly into `world` aligned slices
        cap = max(granule, int(bucket_bytes) // elem_bytes)
        if tail_bytes is None:
            tail_bytes = min(2 << 20, int(bucket_bytes) // 4)
        tail = max(granule, int(tail_bytes) // elem_bytes)
        order = list(reversed(self.params))         # head first == autograd completion order
        sizes = [_round_up(p.numel(), ALIGN_ELEMS) for p in order]
        # walk from the END of the schedule (the stem) to place the two shrinking tail boundaries
        cuts = set()
        acc, i, caps = 0, len(order), [tail, 4 * tail]
        for c in caps:
            if c >= cap:
                break
            acc = 0
            while i > 0 and acc + sizes[i - 1] <= c:
                acc += sizes[i - 1]
                i -= 1
            if i <= 0 or acc == 0:
                break
            cuts.add(i)                             # a bucket starts at parameter i
        self.slots: Dict[int, Slot] = {}
        self.buckets: List[Bucket] = []
        off = 0
        cur: List[torch.nn.Parameter] = []
        cur_start = 0
        for k, p in enumerate(order):
            n = sizes[k]
            if cur and ((off - cur_start) + n > cap or k in cuts):
                end = _round_up(off, granule)
                self.buckets.append(Bucket(len(self.buckets), cur_start, end - cur_start, cur))
                off, cur, cur_start = end, [], end
            self.slots[id(p)] = Slot(off, p.numel(), len(self.buckets))
            cur.append(p)
            off += n
        if cur:
            end = _round_up(off, granule)
            self.buckets.append(Bucket(len(self.buckets), cur_start, end - cur_start, cur))
            off = end
        self.total = off

    def view(self, arena: torch.Tensor, p: torch.nn.Parameter, like: Optional[torch.Tensor] = None) -> torch.Tensor:
        """A view of ``arena`` shaped (and strided) like ``p``.  Channels-last 4-D parameters keep their
        physical KRSC order inside the arena so conv kernels can consume the slot directly."""
        s = self.slots[id(p)]
        flat = arena[s.offset:s.offset + s.numel]
        ref = p if like is None else like
        if ref.dim() == 4 and ref.is_contiguous(memory_format=torch.channels_last) and not ref.is_contiguous():
            k, c, r, q = ref.shape
            return flat.view(k, r, q, c).permute(0, 3, 1, 2)
        return flat.view(ref.shape)

    def bucket_view(self, arena: torch.Tensor, b: Bucket) -> torch.Tensor:
        return arena[b.start:b.start + b.numel]


def alloc_arena(numel: int, dtype: torch.dtype, device: torch.device,
                allocator: Optional[Callable[[int, torch.dtype], torch.Tensor]] = None) -> torch.Tensor:
    numel = max(numel, ALIGN_ELEMS)
    if allocator is not None:
        t = allocator(numel, dtype)
        t.zero_()
        return t
    return torch.zeros(numel, dtype=dtype, device=device)
