"""Training engine (the body of the reference's ``main()``, /root/reference/main.py:49-185).

Flow kept from the reference: build data -> ``initialize_model`` -> Adam(lr=LR) ->
optional resume -> weight sync from rank 0 -> per epoch {train over the shard with
CE loss; rank 0 checkpoints and (optionally) validates} with the same log lines
(``_Epoch: {} | Train Loss: {} | Time: {}``, ``_Epoch: {} | Acc: {}``).

Changed underneath: device-resident bf16 channels-last activations, on-device synthetic
shards (or GPU-preprocessed files), gradients reduced bucket-by-bucket *during*
backward by the fused peer-memory all-reduce+Adam kernel, CUDA-event timing reported as
max over ranks, resume continues from the stored epoch, Inception's aux
loss handled (D2), frozen parameters allowed (D1), logged loss additionally all-reduced
(D9), validation does not stall the other ranks (they wait at a barrier instead of
inside an Allreduce, D8).
"""
from __future__ import annotations

import os
from typing import Any, Dict, List, Optional, Tuple

import torch
import torch.nn as nn

from .. import ops
from ..ops import scratch
from ..data.synthetic import SyntheticShard
from ..models import initialize_model
from ..parallel import bootstrap
from ..parallel.collectives import Hyper
from ..parallel.ddp import FusedDataParallel
from ..utils.config import Settings, image_hw
from ..utils.logging import get_logger
from ..utils.timing import DeviceTimer, max_over_ranks
from .checkpoint import checkpoint_path, load_checkpoint, make_checkpoint, save_checkpoint


def compute_dtype(settings: Settings, device: torch.device) -> torch.dtype:
    if device.type == "cuda" and settings.DTYPE == "bf16":
        return torch.bfloat16
    return torch.float32


def loss_from_outputs(outputs: Any, labels: torch.Tensor) -> torch.Tensor:
    """CE on the main head (+0.4 x CE on Inception's aux head when it exists)."""
    if isinstance(outputs, tuple):
        logits, aux = outputs[0], outputs[1]
        loss = ops.cross_entropy(logits, labels)
        if aux is not None:
            loss = loss + 0.4 * ops.cross_entropy(aux, labels)
        return loss
    return ops.cross_entropy(outputs, labels)


def count_launches(engine) -> int:
    """Kernels of this framework launched so far (ops + tensor-core + peer collectives)."""
    n = int(getattr(engine.coll, "launches", 0))
    if engine.device.type == "cuda" and ops.native_enabled():
        from ..ops import tc
        n += ops.native().LAUNCHES["n"] + tc.LAUNCHES["n"]
    return n


class Trainer:
    """One model replica + its DP engine.  ``train_step`` is the hot loop body of main.py:146-156."""

    def __init__(self, settings: Settings, world: Optional[bootstrap.World] = None, model: Optional[nn.Module] = None):
        self.s = settings
        self.world = world or bootstrap.init_world(settings.DEVICE)
        self.device = self.world.device
        self.dtype = compute_dtype(settings, self.device)
        self.log = get_logger()
        torch.manual_seed(settings.SEED)
        if model is None:
            model, self.input_size = initialize_model(settings.MODEL_NAME, settings.NUM_CLASSES,
                                                      settings.FEATURE_EXTRACT, use_pretrained=settings.USE_PRETRAINED)
        else:
            self.input_size = 224
        self.hw = image_hw(settings, self.input_size)
        self.channels_last = self.device.type == "cuda"
        if self.channels_last:
            model = model.to(memory_format=torch.channels_last)
        self.model = model
        hyper = Hyper(kind=settings.OPTIMIZER, lr=settings.LR, momentum=settings.MOMENTUM)
        shadow = torch.bfloat16 if (self.dtype == torch.bfloat16 and ops.native_enabled()) else None
        wire = torch.bfloat16 if str(settings.GRAD_DTYPE).lower() in ("bf16", "bfloat16") else torch.float32
        self.engine = FusedDataParallel(model, self.world, hyper, bucket_mb=settings.BUCKET_MB,
                                        algo=settings.ALLREDUCE_ALGO, shadow_dtype=shadow, grad_dtype=wire)
        self.optimizer = self.engine.optimizer
        self.start_epoch = 0
        self.autocast = (self.dtype == torch.bfloat16 and not ops.native_enabled())
        self._graph = None
        self._graph_static = None
        self._graph_failed = False
        self.graph_launches_per_step = 0

    # ------------------------------------------------------------------ set-up helpers
    def maybe_resume(self) -> None:
        if self.s.FROM_CHECKPOINT:
            path = os.path.join(self.s.CHECKPOINT_DIR, self.s.CHECKPOINT_NAME)
            self.log.info("_Loading Checkpoint")
            _, _, epoch = load_checkpoint(path, self.model, self.optimizer, map_location=self.device)
            self.engine.refresh_shadow()
            self.start_epoch = epoch + 1
            self.log.info("_Checkpoint loaded")

    def sync_params(self) -> None:
        self.engine.sync_params(0)

    def make_file_shards(self):
        """CSV + JPEG path (SYNTHETIC=False): rank-local training shard and, on rank 0, the un-sharded
        validation set -- the same split logic as main.py:73-112, minus the pickled scatter."""
        from ..data.dataset import FileShard, load_index, render_index_images
        if not os.path.exists(self.s.TEST_CSV if self.s.DEBUG else self.s.TRAIN_CSV):
            raise FileNotFoundError("SYNTHETIC=False needs the CSV index {} (shipped under ./data; or run "
                                    "create_dataset.py)".format(self.s.TEST_CSV if self.s.DEBUG else self.s.TRAIN_CSV))
        train_df, _ = load_index(self.s.TRAIN_CSV, self.s.TEST_CSV, self.s.DEBUG, self.s.SEED)
        if self.world.rank == 0:
            render_index_images(train_df, self.s.TRAIN_DIR)     # stand-in JPEGs for generated rows; no-op for Herbarium
        if self.world.size > 1:
            import torch.distributed as dist
            dist.barrier()
        tr = FileShard(self.s.TRAIN_DIR, train_df, self.s.BATCH_SIZE, self.hw, self.world.rank, self.world.size,
                       self.device, self.dtype, self.channels_last, True, self.s.SEED, mode="train")
        va = None
        if self.s.VALIDATE and self.world.rank == 0:
            va = FileShard(self.s.TRAIN_DIR, train_df, self.s.BATCH_SIZE, self.hw, 0, 1, self.device, self.dtype,
                           self.channels_last, False, self.s.SEED, mode="train")
        return tr, va

    def make_shard(self, total: Optional[int] = None, shuffle: bool = True, all_ranks: bool = True) -> SyntheticShard:
        if total is None:
            total = self.s.SYNTHETIC_SAMPLES or (800 if self.s.DEBUG else 40000)   # main.py:73-82 sizes
        r, w = (self.world.rank, self.world.size) if all_ranks else (0, 1)
        return SyntheticShard(total, self.s.BATCH_SIZE, self.s.NUM_CLASSES, self.hw, r, w, self.s.SEED, self.device,
                              self.dtype, self.channels_last, shuffle)

    # ------------------------------------------------------------------ hot loop
    def forward_loss(self, images: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        if self.autocast:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = self.model(images)
            if isinstance(out, tuple):
                out = tuple(o.float() if o is not None else None for o in out)
            else:
                out = out.float()
            return loss_from_outputs(out, labels)
        return loss_from_outputs(self.model(images), labels)

    def train_step(self, images: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """fwd -> CE -> zero_grad -> backward (buckets reduce+update as they complete) -> join."""
        pool = self._scratch_pool()
        if pool is not None:
            scratch.arm(pool)          # one zeroing launch serves every BN-statistics / BN-backward accumulator
            if self._has_dropout:
                ops.native().advance_dropout_epoch(self.device)     # fresh masks on every CUDA-graph replay
        try:
            loss = self.forward_loss(images, labels)
            self.optimizer.zero_grad()
            loss.backward()
            self.optimizer.step()
        finally:
            if pool is not None:
                scratch.disarm()
        return loss.detach()

    def _scratch_pool(self):
        if self.device.type != "cuda" or not getattr(self.s, "NATIVE_OPS", False):
            return None
        if getattr(self, "_scratch_buf", None) is None:
            self._scratch_buf = scratch.new_buffer(self.device)
        return self._scratch_buf

    # ------------------------------------------------------------------ CUDA-graph replay of the whole step
    def graph_capable(self) -> bool:
        """Whole-step capture needs a CUDA device and fixed shapes.  Dropout is capturable on the native path: its
        kernels mix a device-resident step counter into the seed (``advance_dropout_epoch``); on the torch-ops path
        the masks come from torch's graph-aware CUDA generator."""
        return self.device.type == "cuda"

    @property
    def _has_dropout(self) -> bool:
        if getattr(self, "_has_dropout_cached", None) is None:
            self._has_dropout_cached = any(isinstance(m, nn.Dropout) and m.p > 0 for m in self.model.modules())
        return self._has_dropout_cached

    def train_step_graphed(self, images: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
        """Same work as :meth:`train_step`, but forward + backward + bucketed all-reduce + fused Adam are
        captured once into a CUDA graph and replayed: one host call per step instead of ~600 launches
        (the eager step is host-bound on B200: ~18 ms of Python/launch overhead vs ~17 ms of kernels)."""
        if self._graph is None:
            try:
                self._capture(images, labels)
            except Exception:
                self._graph_failed = True
                raise
        sx, sy, sloss = self._graph_static
        sx.copy_(images, non_blocking=True)
        sy.copy_(labels, non_blocking=True)
        self._graph.replay()
        self.engine.step_count += 1
        return sloss

    def _step_for(self, images: torch.Tensor):
        """The step function ``fit`` uses for a batch of this shape: the CUDA-graph replay for the shape the graph was
        captured with (every full batch), the eager step for anything else (a ragged last batch) or when graphs are
        switched off (``USE_CUDA_GRAPH=False`` / ``MPB_CUDA_GRAPH=0``) or not available (CPU)."""
        if not (self.graph_capable() and getattr(self.s, "USE_CUDA_GRAPH", True)
                and os.environ.get("MPB_CUDA_GRAPH", "1") != "0"):
            return self.train_step
        if self._graph is None:
            if self._graph_failed or images.shape[0] != self.s.BATCH_SIZE:
                return self.train_step
            return self.train_step_graphed      # first full batch: two eager warm-up steps, capture, replay
        return self.train_step_graphed if tuple(images.shape) == tuple(self._graph_static[0].shape) else self.train_step

    def _capture(self, images: torch.Tensor, labels: torch.Tensor) -> None:
        """Warm up (allocator, lazily built tensor maps, autograd graph), then record one step.  The two warm-up steps
        are real optimizer steps, so model / optimizer / BatchNorm state is saved before and put back after them: the
        first replay is then exactly the first training step (``fit`` must not train its first batch three times)."""
        sx, sy = images.clone(), labels.clone()
        eng = self.engine
        with torch.no_grad():
            eng.gather_optimizer_state()
            saved_p = eng.param_arena.clone()
            saved_s = {k: v.clone() for k, v in eng.state.items()}
            saved_b = [b.detach().clone() for b in self.model.buffers()]
            saved_frozen = [p.detach().clone() for p in eng.params_all if not p.requires_grad]
            saved_step = eng.step_count
        cur = torch.cuda.current_stream(self.device)
        side = torch.cuda.Stream(self.device)
        side.wait_stream(cur)
        with torch.cuda.stream(side):
            for _ in range(2):                       # warm-up on a side stream
                self.train_step(sx, sy)
        cur.wait_stream(side)
        with torch.no_grad():
            eng.param_arena.copy_(saved_p)
            for k, v in eng.state.items():
                v.copy_(saved_s[k])
            for b, sb in zip(self.model.buffers(), saved_b):
                b.copy_(sb)
            for p, sp in zip([p for p in eng.params_all if not p.requires_grad], saved_frozen):
                p.copy_(sp)
            eng.set_step(saved_step)
            eng.refresh_shadow()
        torch.cuda.synchronize(self.device)
        if self.world.size > 1:
            import torch.distributed as dist
            dist.barrier()                           # every rank restored before anyone's replay writes peers
        g = torch.cuda.CUDAGraph()
        n0 = count_launches(self.engine)
        with torch.cuda.graph(g):
            sloss = self.train_step(sx, sy)
        self.graph_launches_per_step = count_launches(self.engine) - n0     # our kernels inside one replay
        self.engine.step_count = saved_step          # capture does not execute: no step was taken
        self._graph, self._graph_static = g, (sx, sy, sloss)

    @torch.no_grad()
    def evaluate(self, shard: SyntheticShard) -> float:
        """Rank-local accuracy pass (main.py:173-185): argmax, compare, count."""
        self.model.eval()
        correct = torch.zeros((), dtype=torch.int64, device=self.device)
        n = 0
        for images, labels in shard:
            if self.autocast:
                with torch.autocast("cuda", dtype=torch.bfloat16):
                    out = self.model(images)
            else:
                out = self.model(images)
            correct += ops.count_correct(out, labels)
            n += labels.numel()
        self.model.train()
        return float(correct.item()) / max(n, 1)

    # ------------------------------------------------------------------ the epoch loop
    def fit(self, train: Optional[SyntheticShard] = None, val: Optional[SyntheticShard] = None,
            epochs: Optional[int] = None) -> List[Dict[str, float]]:
        s, log, w = self.s, self.log, self.world
        if train is None and not s.SYNTHETIC:
            train, val = self.make_file_shards()
        train = train or self.make_shard()
        log.info("_Files Received: {}".format(train.num_samples))
        if val is None and s.VALIDATE and w.rank == 0:
            val = self.make_shard(total=train.total, shuffle=False, all_ranks=False)   # the un-sharded sample, main.py:108
        log.info("_Model Created: {}".format(s.MODEL_NAME))
        self.maybe_resume()
        self.sync_params()
        log.info("_Model loaded to {}".format(str(self.device).upper()))
        log.info("_Entering training Loop")
        history: List[Dict[str, float]] = []
        n_epochs = s.NUM_EPOCHS if epochs is None else epochs
        for epoch in range(self.start_epoch, n_epochs):
            self.model.train()
            train.set_epoch(epoch)
            tr_loss = torch.zeros((), dtype=torch.float32, device=self.device)
            with DeviceTimer(self.device) as t:
                for images, labels in train:
                    tr_loss += self._step_for(images)(images, labels).float()
            nb = max(len(train), 1)
            local = float(tr_loss.item()) / nb
            secs = max_over_ranks(t.ms, self.device if w.backend == "nccl" else "cpu") / 1e3
            log.info("_Epoch: {} | Train Loss: {} | Time: {}".format(epoch, local, secs))
            rec = {"epoch": epoch, "loss": local, "time_s": secs,
                   "img_per_s": train.num_samples * w.size / max(secs, 1e-9)}
            if w.rank == 0:
                log.info("_Creating a checkpoint at epoch {}".format(epoch))
            ck = make_checkpoint(epoch, self.model, self.optimizer, local)       # collective (gathers sliced state)
            if w.rank == 0:
                save_checkpoint(ck, epoch, s.MODEL_NAME, s.CHECKPOINT_DIR, s.MODELS_DIR)
                log.info("_Checkpoint saved")
                if s.VALIDATE and val is not None:
                    log.info("_Evaluating model")
                    rec["acc"] = self.evaluate(val)
                    log.info("_Epoch: {} | Acc: {}".format(epoch, rec["acc"]))
            if w.size > 1:
                import torch.distributed as dist
                dist.barrier()
            history.append(rec)
        return history

    def close(self) -> None:
        self.engine.close()


def run_training(settings: Settings, world: Optional[bootstrap.World] = None) -> List[Dict[str, float]]:
    tr = Trainer(settings, world)
    try:
        return tr.fit()
    finally:
        tr.close()
