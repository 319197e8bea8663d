"""Synchronous data parallelism: bucketed, overlapped, optimizer-fused gradient all-reduce.

What the reference does per step (/root/reference/main.py:151-155,
mpi_tools.py:30-37): ``zero_grad`` -> ``backward`` -> *after backward has finished*, one
blocking ``Allreduce`` + host divide per parameter tensor -> ``optimizer.step()``.

What this engine does: parameters, gradients and Adam moments live in flat arenas
(parallel/arena.py).  ``p.grad`` is a permanent view into the gradient arena, so
autograd accumulates straight into the communication buffer.  A
post-accumulate-grad hook counts down each bucket; the moment a bucket is complete its
fused kernel (sum over peers x 1/world -> Adam/SGD update -> write back fp32 params and
the bf16 compute copy) is enqueued on a dedicated communication stream while autograd
keeps producing the next bucket.  ``step()`` is then just "wait for the comm stream".
"""
from __future__ import annotations

from typing import Any, Dict, Iterable, List, Optional

import torch
import torch.nn as nn

from .arena import ALIGN_ELEMS, Bucket, FlatLayout, alloc_arena
from .collectives import Hyper, make_collectives


class FusedOptimizer(torch.optim.Optimizer):
    """Optimizer facade whose state tensors are views into the engine's flat arenas.

    ``state_dict()`` (inherited) therefore emits exactly ``torch.optim.Adam``'s layout --
    ``state[idx] = {step, exp_avg, exp_avg_sq}`` and one ``param_groups`` entry listing
    parameter indices in ``model.parameters()`` order -- which is what the reference's
    ``helpers.load_checkpoint`` feeds to a stock Adam (helpers.py:14)."""

    def __init__(self, engine: "FusedDataParallel", params: List[nn.Parameter], hyper: Hyper):
        if hyper.kind == "adam":
            defaults = dict(lr=hyper.lr, betas=(hyper.beta1, hyper.beta2), eps=hyper.eps,
                            weight_decay=hyper.weight_decay, amsgrad=False, maximize=False, foreach=None,
                            capturable=False, differentiable=False, fused=None, decoupled_weight_decay=False)
        else:
            defaults = dict(lr=hyper.lr, momentum=hyper.momentum, dampening=0, weight_decay=hyper.weight_decay,
                            nesterov=False, maximize=False, foreach=None, differentiable=False, fused=None)
        super().__init__(params, defaults)
        self._engine = engine
        self._hyper = hyper

    def _bind_state(self) -> None:
        eng = self._engine
        for p in eng.layout.params:
            st = self.state[p]
            if self._hyper.kind == "adam":
                st["step"] = torch.tensor(float(eng.step_count), dtype=torch.float32)
                st["exp_avg"] = eng.layout.view(eng.state["exp_avg"], p)
                st["exp_avg_sq"] = eng.layout.view(eng.state["exp_avg_sq"], p)
            else:
                st["momentum_buffer"] = eng.layout.view(eng.state["momentum_buffer"], p)

    def state_dict(self) -> Dict[str, Any]:
        self._engine.gather_optimizer_state()
        self._bind_state()
        return super().state_dict()

    def load_state_dict(self, state_dict: Dict[str, Any]) -> None:
        eng = self._engine
        groups = state_dict["param_groups"]
        ids = [i for g in groups for i in g["params"]]
        params = [p for g in self.param_groups for p in g["params"]]
        if len(ids) != len(params):
            raise ValueError("optimizer state has {} parameters, model has {}".format(len(ids), len(params)))
        step = 0
        for idx, p in zip(ids, params):
            st = state_dict["state"].get(idx)
            if st is None or id(p) not in eng.layout.slots:
                continue
            for key in eng.state:
                if key in st:
                    eng.layout.view(eng.state[key], p).copy_(st[key].to(eng.state[key].dtype).reshape(p.shape))
            if "step" in st:
                step = max(step, int(float(st["step"])))
        eng.set_step(step)
        g0 = groups[0]
        self.param_groups[0]["lr"] = g0.get("lr", self.param_groups[0]["lr"])
        self._hyper.lr = float(self.param_groups[0]["lr"])
        self._bind_state()
        eng.refresh_shadow()          # a resume loads the model weights just before this call

    def zero_grad(self, set_to_none: bool = False) -> None:      # grads are arena views: never drop them
        self._engine.zero_grad()

    @torch.no_grad()
    def step(self, closure=None):
        self._hyper.lr = float(self.param_groups[0]["lr"])
        self._engine.finish_step()
        return None


class FusedDataParallel:
    """Owns the arenas, the hooks and the communication stream for one model replica."""

    def __init__(self, model: nn.Module, world, hyper: Optional[Hyper] = None, bucket_mb: float = 32.0,
                 algo: str = "auto", grad_dtype: torch.dtype = torch.float32,
                 shadow_dtype: Optional[torch.dtype] = None, overlap: bool = True, collectives=None,
                 tail_bytes: Optional[int] = None):
        self.model, self.world = model, world
        self.device = world.device
        self.hyper = hyper or Hyper()
        self.algo = algo
        self.overlap = overlap and self.device.type == "cuda"
        self.coll = collectives if collectives is not None else make_collectives(world, algo)
        model.to(self.device)
        self.params_all = list(model.parameters())
        self.layout = FlatLayout(self.params_all, int(bucket_mb * 2 ** 20), world.size, tail_bytes=tail_bytes)
        n = self.layout.total
        dev = self.device
        self.param_arena = alloc_arena(n, torch.float32, dev, self.coll.alloc)
        # Gradients are always ACCUMULATED in fp32 (split-K wgrad adds partial sums into this arena).  With
        # grad_dtype=bf16 the exchange itself is bf16: the fused kernel packs each bucket into `wire_arena`, peers
        # read 2 bytes per element, sums are fp32 and the optimizer works on the fp32 masters.
        self.grad_arena = alloc_arena(n, torch.float32, dev, self.coll.alloc)
        self.wire_dtype = grad_dtype
        self.wire_arena = alloc_arena(n, torch.bfloat16, dev, self.coll.alloc) \
            if (grad_dtype == torch.bfloat16 and world.size > 1) else None
        self.shadow_arena = alloc_arena(n, shadow_dtype, dev, self.coll.alloc) if shadow_dtype is not None else None
        keys = ("exp_avg", "exp_avg_sq") if self.hyper.kind == "adam" else ("momentum_buffer",)
        self.state = {k: torch.zeros(max(n, 1), dtype=torch.float32, device=dev) for k in keys}
        self.step_count = 0
        # device-resident mirror of the optimizer step: the fused kernels read it, so a captured CUDA graph
        # keeps applying the right Adam bias corrections on every replay
        self.step_dev = torch.zeros(1, dtype=torch.int32, device=dev) if dev.type == "cuda" else None
        self._step_advanced = False
        if hasattr(self.coll, "step_dev"):
            self.coll.step_dev = self.step_dev
        self._rebind_parameters()
        for b in self.layout.buckets:
            b.algo = algo
        self._hooks = []
        self._in_flight: List[Bucket] = []
        self._state_sliced = False
        self._suspended = False
        if self.device.type == "cuda":
            self.comm_stream = torch.cuda.Stream(device=dev)
        else:
            self.comm_stream = None
        for p in self.layout.params:
            self._hooks.append(p.register_post_accumulate_grad_hook(self._on_grad))
        # any in-place weight load (checkpoint resume, user code) must also refresh the bf16 compute copy
        self._hooks.append(model.register_load_state_dict_post_hook(lambda _m, _keys: self.refresh_shadow()))
        self.optimizer = FusedOptimizer(self, self.layout.params, self.hyper)
        self.optimizer._bind_state()
        self._reset_pending()

    # ------------------------------------------------------------------ set-up
    @torch.no_grad()
    def _rebind_parameters(self) -> None:
        """Move every trainable parameter's storage into the arena; p.grad becomes an arena view."""
        for p in self.layout.params:
            v = self.layout.view(self.param_arena, p)
            v.copy_(p.data)
            p.data = v
            p.grad = self.layout.view(self.grad_arena, p)
        if self.shadow_arena is not None:
            self.shadow_arena.copy_(self.param_arena)
            for p in self.layout.params:
                p._mpb_shadow = self.layout.view(self.shadow_arena, p)      # bf16 compute copy read by ops.cuda_impl
        if self.device.type == "cuda":
            for p in self.layout.params:
                p._mpb_grad_ready = self._on_grad     # kernels that write p.grad in place report completion here

    def _reset_pending(self) -> None:
        for b in self.layout.buckets:
            b.pending = len(b.params)
        self._fired = set()
        self._in_flight = []

    # ------------------------------------------------------------------ weight sync (C14 / K2)
    @torch.no_grad()
    def sync_params(self, root: int = 0) -> None:
        """Broadcast rank-``root`` parameters *and buffers* (the reference forgets BN running
        statistics) -- one flat broadcast instead of one Bcast per tensor
        (mpi_tools.py:47-53)."""
        if self.world.size == 1:
            self.refresh_shadow()
            return
        self.coll.broadcast_(self.param_arena[: self.layout.total], root)
        self.refresh_shadow()
        frozen = [p for p in self.params_all if not p.requires_grad]
        extras = [t for t in list(frozen) + list(self.model.buffers()) if t.numel() > 0]
        if extras:
            flat = torch.cat([t.detach().reshape(-1).to(torch.float32) for t in extras])
            # the peer broadcast moves 16-byte vectors: stage a zero-tailed multiple of ALIGN_ELEMS floats
            # (DenseNet-121's buffers total 83,769 elements)
            n_pad = -(-flat.numel() // ALIGN_ELEMS) * ALIGN_ELEMS
            stage = self.coll.alloc(n_pad, torch.float32)
            stage.zero_()
            stage[:flat.numel()].copy_(flat)
            self.coll.broadcast_(stage, root)
            off = 0
            for t in extras:
                k = t.numel()
                t.detach().copy_(stage[off:off + k].view(t.shape).to(t.dtype))
                off += k
        self.coll.sync_stream()

    @torch.no_grad()
    def refresh_shadow(self) -> None:
        """Re-derive the bf16 compute copy from the fp32 master arena.  Needed after anything that writes
        parameters behind the fused kernels' back: ``load_state_dict`` / checkpoint resume, manual edits."""
        if self.shadow_arena is not None:
            self.shadow_arena.copy_(self.param_arena)

    # ------------------------------------------------------------------ per-step machinery
    def zero_grad(self) -> None:
        self.grad_arena.zero_()
        self._reset_pending()
        if self.step_dev is not None and not self._step_advanced:
            self.step_dev.add_(1)                 # once per optimizer step, before any bucket launches
            self._step_advanced = True

    def _launch(self, b: Bucket) -> None:
        g = self.layout.bucket_view(self.grad_arena, b)
        p = self.layout.bucket_view(self.param_arena, b)
        st = {k: self.layout.bucket_view(v, b) for k, v in self.state.items()}
        sh = self.layout.bucket_view(self.shadow_arena, b) if self.shadow_arena is not None else None
        kw = {"wire": self.layout.bucket_view(self.wire_arena, b)} if self.wire_arena is not None else {}
        step = self.step_count + 1
        if self.comm_stream is not None and self.overlap:
            self.comm_stream.wait_stream(torch.cuda.current_stream(self.device))
            with torch.cuda.stream(self.comm_stream):
                self.coll.all_reduce_update_(g, p, st, sh, self.hyper, step, b.algo, **kw)
        else:
            self.coll.all_reduce_update_(g, p, st, sh, self.hyper, step, b.algo, **kw)
        if not self.coll.owns_full_state(b.algo):
            self._state_sliced = True
        self._in_flight.append(b)

    def suspended(self):
        """Context manager: run forward/backward with the engine detached -- gradients of THIS rank's batch stay in
        the gradient arena, nothing is exchanged, no optimizer step is taken.  Used to check the fused path against
        a plain all-reduce (bench.py ``dp_check``) and for gradient inspection."""
        import contextlib

        @contextlib.contextmanager
        def _cm():
            self._suspended = True
            try:
                yield self
            finally:
                self._suspended = False
                self._reset_pending()
        return _cm()

    def _on_grad(self, p: nn.Parameter) -> None:
        if self._suspended:
            return
        # Idempotent per step: a kernel that wrote p.grad in place reports completion explicitly, and autograd may
        # additionally run the post-accumulate hook for the same parameter (observed on B200 / torch 2.11).
        if id(p) in self._fired:
            return
        self._fired.add(id(p))
        b = self.layout.buckets[self.layout.slots[id(p)].bucket]
        b.pending -= 1
        if b.pending == 0:
            with torch.no_grad():
                self._launch(b)

    @torch.no_grad()
    def finish_step(self) -> None:
        """Flush buckets whose hooks did not all fire (unused parameters) and join the comm stream."""
        for b in self.layout.buckets:
            if b.pending > 0 and b not in self._in_flight:
                self._launch(b)
        if self.comm_stream is not None and self.overlap:
            torch.cuda.current_stream(self.device).wait_stream(self.comm_stream)
        self.step_count += 1
        self._step_advanced = False
        self._reset_pending()

    @torch.no_grad()
    def reduce_and_step(self) -> None:
        """Exchange + optimizer step over whatever the gradient arena holds NOW (every bucket, in schedule order).
        With :meth:`suspended` this splits a step into "compute local gradients" and "reduce + update", which is
        what lets a check feed the fused kernels and a plain all-reduce the very same gradients."""
        if self.step_dev is not None and not self._step_advanced:
            self.step_dev.add_(1)
            self._step_advanced = True
        self._reset_pending()
        self.finish_step()

    def set_step(self, step: int) -> None:
        self.step_count = int(step)
        if self.step_dev is not None:
            self.step_dev.fill_(int(step))

    @torch.no_grad()
    def average_gradients_only(self) -> None:
        """``mpi_avg_grads`` semantics: gradients averaged in place, no optimizer update."""
        if self.world.size == 1:
            return
        for b in self.layout.buckets:
            self.coll.all_reduce_avg_(self.layout.bucket_view(self.grad_arena, b), b.algo)
        self.coll.sync_stream()

    @torch.no_grad()
    def gather_optimizer_state(self) -> None:
        """Two-shot buckets update only the owner's slice of the moments (ZeRO-1 data flow);
        make the full arrays valid before a checkpoint."""
        if self._state_sliced and self.world.size > 1:
            for b in self.layout.buckets:
                if not self.coll.owns_full_state(b.algo):
                    for v in self.state.values():
                        self.coll.all_gather_slices_(self.layout.bucket_view(v, b))
            self.coll.sync_stream()

    def close(self) -> None:
        for h in self._hooks:
            h.remove()
        self._hooks = []
        self.coll.close()
