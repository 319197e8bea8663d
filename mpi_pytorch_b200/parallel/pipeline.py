"""Rank-staged evaluation pipeline: read -> resize -> normalise+route -> predict.

Role map of the reference (/root/reference/evaluation_pipeline.py:162-199): rank 0 reads
images and assigns every image a predictor rank uniformly at random, rank 1 resizes,
rank 2 tensorises + normalises and routes each image to *its* predictor, ranks 3..N-1
hold a model replica each and count correct predictions; the partial results are summed
on rank 0.  The reference moves one pickled PIL image per blocking ``comm.send``.

Here every edge carries **fixed-shape device tensors** -- a small int64 header plus a
micro-batch payload -- over ``torch.distributed`` p2p (NCCL ``ncclSend/ncclRecv`` on
GPUs, gloo on CPU), double-buffered with ``isend`` so a stage's kernel overlaps the
previous hand-off.  Resize and normalise are CUDA kernels (``ops.resize_normalize`` and
friends).  Per-image routing, the ``None``-sentinel termination protocol (a header with
``last=1``) and the "sum of partial results" reduction are preserved; the reduction is an
exact int64 count instead of a float.  With fewer than 4 ranks the stages fold onto the
available ranks instead of raising (``np.random.randint(3, size)`` needs size > 3 in the
reference, evaluation_pipeline.py:178).
"""
from __future__ import annotations

import dataclasses
from typing import Callable, Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

from .. import ops
from ..utils.logging import get_logger

HDR_ROWS_EXTRA = 1        # row 0 of every header: (count, last, reserved)


@dataclasses.dataclass
class Roles:
    reader: int
    resizer: int
    normalizer: int
    predictors: List[int]


def assign_roles(size: int) -> Roles:
    if size >= 4:
        return Roles(0, 1, 2, list(range(3, size)))
    if size == 3:
        return Roles(0, 0, 1, [2])
    if size == 2:
        return Roles(0, 0, 0, [1])
    return Roles(0, 0, 0, [0])


class _Edge:
    """One directed p2p edge with double-buffered asynchronous sends."""

    def __init__(self, me: int, peer: int, device: torch.device):
        self.me, self.peer, self.device = me, peer, device
        self._pending: List[Tuple[object, torch.Tensor]] = []

    def send(self, header: torch.Tensor, payload: Optional[torch.Tensor]) -> None:
        works = [(dist.isend(header, dst=self.peer), header)]
        if payload is not None:
            works.append((dist.isend(payload, dst=self.peer), payload))
        self._pending.extend(works)
        while len(self._pending) > 4:              # at most two messages (hdr+payload) in flight
            w, _keep = self._pending.pop(0)
            w.wait()

    def flush(self) -> None:
        for w, _keep in self._pending:
            w.wait()
        self._pending = []

    def recv(self, header: torch.Tensor, payload: torch.Tensor) -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
        dist.recv(header, src=self.peer)
        if int(header[0, 0].item()) > 0:
            dist.recv(payload, src=self.peer)
            return header, payload
        return header, None


def make_header(mb: int, device: torch.device) -> torch.Tensor:
    return torch.zeros(mb + HDR_ROWS_EXTRA, 3, dtype=torch.int64, device=device)


class EvalPipeline:
    """Drives one rank's part of the pipeline.

    ``source(i0, i1) -> uint8 [n, RAW_H, RAW_W, 3]`` yields raw frames for global sample ids
    ``[i0, i1)``; ``labels`` is the int64 label vector of the whole test set (rank 0 only
    needs it); ``model_fn()`` builds the predictor's model (with the checkpoint loaded)."""

    def __init__(self, world, n_samples: int, raw_hw: Tuple[int, int], out_hw: Tuple[int, int],
                 source: Callable[[int, int], torch.Tensor], labels: torch.Tensor,
                 model_fn: Callable[[], torch.nn.Module], micro_batch: int = 32, seed: int = 0,
                 dtype: torch.dtype = torch.float32, resize_mode: str = "bicubic", collectives=None):
        self.w = world
        self.coll = collectives        # optional: sums the per-rank counts with the peer-memory count kernel (K13)
        self.dev = world.device
        self.roles = assign_roles(world.size)
        self.n, self.mb, self.seed = int(n_samples), int(micro_batch), seed
        self.raw_hw, self.out_hw = raw_hw, out_hw
        self.source, self.labels, self.model_fn = source, labels, model_fn
        self.dtype, self.resize_mode = dtype, resize_mode
        self.log = get_logger()
        self.correct = torch.zeros((), dtype=torch.int64, device=self.dev)
        self.seen = 0
        self._model: Optional[torch.nn.Module] = None
        self._edges: Dict[int, _Edge] = {}
        self._queues: Dict[int, List[Tuple[torch.Tensor, int]]] = {}

    # ------------------------------------------------------------------ plumbing
    def _edge(self, peer: int) -> _Edge:
        if peer not in self._edges:
            self._edges[peer] = _Edge(self.w.rank, peer, self.dev)
        return self._edges[peer]

    def _emit(self, dest: int, stage: str, header: torch.Tensor, payload: Optional[torch.Tensor]) -> None:
        if dest == self.w.rank:
            getattr(self, "_stage_" + stage)(header, payload)
        else:
            self._edge(dest).send(header, payload)

    # ------------------------------------------------------------------ stage 0: read + dispatch (C19)
    def _run_reader(self) -> None:
        g = torch.Generator().manual_seed(self.seed)
        preds = self.roles.predictors
        assign = torch.tensor(preds)[torch.randint(0, len(preds), (self.n,), generator=g)]   # seeded, unlike :178
        for i0 in range(0, self.n, self.mb):
            i1 = min(i0 + self.mb, self.n)
            k = i1 - i0
            raw = self.source(i0, i1).to(self.dev)
            if k < self.mb:
                pad = torch.zeros(self.mb - k, *raw.shape[1:], dtype=raw.dtype, device=self.dev)
                raw = torch.cat([raw, pad])
            hdr = make_header(self.mb, self.dev)
            hdr[0, 0] = k
            hdr[1:k + 1, 0] = self.labels[i0:i1].to(self.dev)
            hdr[1:k + 1, 1] = assign[i0:i1].to(self.dev)
            hdr[1:k + 1, 2] = torch.arange(i0, i1, device=self.dev)
            self._emit(self.roles.resizer, "resize", hdr, raw)
        end = make_header(self.mb, self.dev)
        end[0, 1] = 1
        self._emit(self.roles.resizer, "resize", end, None)

    # ------------------------------------------------------------------ stage 1: resize (C20)
    def _stage_resize(self, hdr: torch.Tensor, raw: Optional[torch.Tensor]) -> None:
        out = None
        if raw is not None:
            out = ops.resize_u8(raw, self.out_hw[0], self.out_hw[1], self.resize_mode)
        self._emit(self.roles.normalizer, "normalize", hdr, out)

    # ------------------------------------------------------------------ stage 2: normalise + route (C21)
    def _stage_normalize(self, hdr: torch.Tensor, small: Optional[torch.Tensor]) -> None:
        last = int(hdr[0, 1].item()) == 1
        if small is not None:
            k = int(hdr[0, 0].item())
            x = ops.normalize_u8(small, self.dtype)                   # [mb, 3, h, w]
            rows = hdr[1:k + 1].tolist()
            for j, (label, dest, _sid) in enumerate(rows):
                q = self._queues.setdefault(int(dest), [])
                q.append((x[j], int(label)))
                if len(q) == self.mb:
                    self._flush_queue(int(dest))
        if last:
            for dest in self.roles.predictors:
                if self._queues.get(dest):
                    self._flush_queue(dest)
                end = make_header(self.mb, self.dev)
                end[0, 1] = 1
                self._emit(dest, "predict", end, None)

    def _flush_queue(self, dest: int) -> None:
        q = self._queues[dest]
        k = len(q)
        batch = torch.zeros(self.mb, 3, self.out_hw[0], self.out_hw[1], dtype=self.dtype, device=self.dev)
        batch[:k] = torch.stack([t for t, _ in q])
        hdr = make_header(self.mb, self.dev)
        hdr[0, 0] = k
        hdr[1:k + 1, 0] = torch.tensor([l for _, l in q], dtype=torch.int64, device=self.dev)
        self._queues[dest] = []
        self._emit(dest, "predict", hdr, batch)

    # ------------------------------------------------------------------ stage 3: predict (C22)
    def _stage_predict(self, hdr: torch.Tensor, batch: Optional[torch.Tensor]) -> None:
        if batch is None:
            return
        if self._model is None:
            self._model = self.model_fn().to(self.dev).eval()
        k = int(hdr[0, 0].item())
        with torch.no_grad():
            x = batch[:k]
            if self.dev.type == "cuda":
                x = x.contiguous(memory_format=torch.channels_last)
            out = self._model(x)
        self.correct += ops.count_correct(out.float(), hdr[1:k + 1, 0])
        self.seen += k

    # ------------------------------------------------------------------ receive loops
    def _recv_loop(self, src: int, stage: str, payload_shape: Tuple[int, ...], payload_dtype: torch.dtype) -> None:
        edge = self._edge(src)
        while True:
            hdr = make_header(self.mb, self.dev)
            payload = torch.empty(payload_shape, dtype=payload_dtype, device=self.dev)
            hdr, payload = edge.recv(hdr, payload)
            getattr(self, "_stage_" + stage)(hdr, payload)
            if int(hdr[0, 1].item()) == 1:
                break

    def run(self) -> Optional[float]:
        """Execute this rank's role(s); returns the global accuracy on rank 0 (``None`` elsewhere)."""
        r, roles = self.w.rank, self.roles
        raw_shape = (self.mb, self.raw_hw[0], self.raw_hw[1], 3)
        small_shape = (self.mb, self.out_hw[0], self.out_hw[1], 3)
        batch_shape = (self.mb, 3, self.out_hw[0], self.out_hw[1])
        if r == roles.reader:
            self._run_reader()
        elif r == roles.resizer:
            self._recv_loop(roles.reader, "resize", raw_shape, torch.uint8)
        elif r == roles.normalizer:
            self._recv_loop(roles.resizer, "normalize", small_shape, torch.uint8)
        elif r in roles.predictors:
            if self._model is None:
                self._model = self.model_fn().to(self.dev).eval()
            self._recv_loop(roles.normalizer, "predict", batch_shape, self.dtype)
        for e in self._edges.values():
            e.flush()
        if r in roles.predictors:
            self.log.info("Finished node {}, acc {}".format(r, float(self.correct.item()) / max(self.n, 1)))
        total = self.correct.clone()
        if self.w.size > 1:
            if self.coll is not None:
                self.coll.all_reduce_sum_i64(total)     # exact int64 count through peer memory (C23 / K13)
            else:
                dist.all_reduce(total)
        acc = float(total.item()) / max(self.n, 1)
        if r == 0:
            self.log.info("Accuracy is {}".format(acc))
            return acc
        return None
