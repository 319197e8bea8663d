"""On-device synthetic data shards (replaces CSV + ``comm.scatter`` + DataLoader).

The reference's rank 0 reads a CSV, ``np.array_split``s it and pickles the pieces to the
other ranks (/root/reference/main.py:73-91), then every rank decodes JPEGs on the
critical path (data_loader.py:29-37).  BASELINE.json prescribes synthetic data, so here
*nothing* is communicated: sample ``i`` of the global dataset is a pure function of
``(seed, i)`` (a counter-based integer hash), each rank owns the strided index set
``{i : i % world == rank}`` truncated to equal length (no uneven-shard hang),
and batches are produced directly in device memory (CUDA kernel K14 via
``ops.native().synth_batch``; an identical integer-exact torch formulation on CPU).
"""
from __future__ import annotations

from typing import Iterator, Optional, Tuple

import torch

_M32 = 0xFFFFFFFF


def _hash32(x: torch.Tensor) -> torch.Tensor:
    """murmur3 finaliser on int64 tensors holding uint32 values (wrap-around multiply)."""
    x = x & _M32
    x = x ^ (x >> 16)
    x = (x * 0x85EBCA6B) & _M32
    x = x ^ (x >> 13)
    x = (x * 0xC2B2AE35) & _M32
    x = x ^ (x >> 16)
    return x


def synth_labels(seed: int, index: torch.Tensor, num_classes: int) -> torch.Tensor:
    h = _hash32(index.to(torch.int64) * 0x9E3779B1 + (seed * 0x7F4A7C15 + 0x1234567))
    return (h % num_classes).to(torch.int64)


def synth_images_torch(seed: int, index: torch.Tensor, c: int, h: int, w: int,
                       dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """[len(index), c, h, w] values in [-1.7, 1.7) (unit-ish variance).  Element ``e`` of
    sample ``i`` is hash(seed, i*CHW + e) with e in NCHW order; the CUDA kernel computes
    the same number for the same logical coordinate whatever the memory format."""
    n = index.numel()
    chw = c * h * w
    e = torch.arange(chw, dtype=torch.int64, device=index.device).view(1, chw)
    ctr = index.to(torch.int64).view(n, 1) * chw + e
    hsh = _hash32(ctr ^ ((seed * 0x9E3779B9) & _M32))
    u = hsh.to(torch.float64) * (1.0 / 4294967296.0)
    return ((u * 2.0 - 1.0) * 1.7).to(dtype).view(n, c, h, w)


class SyntheticShard:
    """This rank's slice of a synthetic dataset of ``total`` samples."""

    def __init__(self, total: int, batch_size: int, num_classes: int, hw: Tuple[int, int], rank: int = 0,
                 world: int = 1, seed: int = 0, device: torch.device | str = "cpu",
                 dtype: torch.dtype = torch.float32, channels_last: bool = False, shuffle: bool = True,
                 drop_last: bool = False):
        self.total, self.batch_size, self.num_classes = int(total), int(batch_size), int(num_classes)
        self.h, self.w = int(hw[0]), int(hw[1])
        self.rank, self.world, self.seed = rank, world, int(seed)
        self.device = torch.device(device)
        self.dtype, self.channels_last, self.shuffle, self.drop_last = dtype, channels_last, shuffle, drop_last
        self.per_rank = self.total // world            # equal on every rank by construction
        self.indices = torch.arange(self.per_rank, dtype=torch.int64) * world + rank
        self.epoch = 0

    def __len__(self) -> int:
        if self.drop_last:
            return self.per_rank // self.batch_size
        return (self.per_rank + self.batch_size - 1) // self.batch_size

    @property
    def num_samples(self) -> int:
        return self.per_rank

    def set_epoch(self, epoch: int) -> None:
        self.epoch = int(epoch)

    def make_batch(self, idx: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        idx_dev = idx.to(self.device)
        labels = synth_labels(self.seed, idx_dev, self.num_classes)
        if self.device.type == "cuda":
            from .. import ops
            if ops.native_enabled():
                img = ops.native().synth_batch(self.seed, idx_dev, 3, self.h, self.w, self.dtype, self.channels_last)
                return img, labels
        img = synth_images_torch(self.seed, idx_dev, 3, self.h, self.w, self.dtype)
        if self.channels_last:
            img = img.contiguous(memory_format=torch.channels_last)
        return img, labels

    def __iter__(self) -> Iterator[Tuple[torch.Tensor, torch.Tensor]]:
        order = self.indices
        if self.shuffle:
            g = torch.Generator().manual_seed(self.seed * 1000003 + self.epoch * 7919 + self.rank)
            order = order[torch.randperm(self.per_rank, generator=g)]
        nb = len(self)
        for b in range(nb):
            yield self.make_batch(order[b * self.batch_size:(b + 1) * self.batch_size])


class HostStagedShard:
    """End-to-end variant: batches live in *pinned host memory* and are copied H2D every
    step on a side stream (what a real input pipeline does; used by bench.py's e2e number)."""

    def __init__(self, shard: SyntheticShard, n_host_batches: int = 4, device: Optional[torch.device] = None):
        self.device = torch.device(device) if device is not None else shard.device
        cpu = SyntheticShard(shard.total, shard.batch_size, shard.num_classes, (shard.h, shard.w), shard.rank,
                             shard.world, shard.seed, "cpu", shard.dtype, shard.channels_last, False, True)
        self.host = []
        it = iter(cpu)
        for _ in range(max(1, min(n_host_batches, len(cpu)))):
            img, lab = next(it)
            if torch.cuda.is_available():
                img, lab = img.pin_memory(), lab.pin_memory()
            self.host.append((img, lab))
        self.bytes_per_step = self.host[0][0].numel() * self.host[0][0].element_size() + \
            self.host[0][1].numel() * self.host[0][1].element_size()

    def fetch(self, step: int) -> Tuple[torch.Tensor, torch.Tensor]:
        img, lab = self.host[step % len(self.host)]
        return img.to(self.device, non_blocking=True), lab.to(self.device, non_blocking=True)
