#!/usr/bin/env python
"""Headline benchmark: ResNet-34 data-parallel training throughput (images/s), bf16, synthetic
ImageNet-shaped batches of 256 per GPU (BASELINE.json config #2), on N GPUs of one node.

  python bench.py [--gpus 1] [--steps K] [--warmup W]                # this framework
  torchrun --nproc-per-node N ... bench.py --gpus N --steps K --warmup W
  python bench.py --impl reference ...      # the unmodified reference (CPU, mpi4py shim), /root/reference/main.py:146-156
  python bench.py --impl ddp ...            # baseline (b): stock torchvision + torch DDP/NCCL + cuDNN, eager AND graphed
  python bench.py --model vgg16|densenet|inception|...               # BASELINE.json configs #3 / #4
  torchrun --nproc-per-node 8 ... bench.py --task pipeline           # BASELINE.json config #5 (3 + 5 ranks)

Prints ONE JSON line on rank 0 (contract in the task description): ``value`` is the whole-job
device-timed img/s (CUDA events around exactly K steps, max over ranks); ``e2e`` is the same metric
through the public Trainer API including a pinned-host -> device copy of every batch and a device ->
host read of every step's loss.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.abspath(__file__))


def parse_args(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "ddp"])
    ap.add_argument("--task", default="train", choices=["train", "pipeline"])
    ap.add_argument("--model", default=os.environ.get("MPB_BENCH_MODEL", "resnet34"))
    ap.add_argument("--batch", type=int, default=int(os.environ.get("MPB_BENCH_BATCH", "256")))
    ap.add_argument("--image", type=int, default=int(os.environ.get("MPB_BENCH_IMAGE", "0")))
    ap.add_argument("--classes", type=int, default=int(os.environ.get("MPB_BENCH_CLASSES", "1000")))
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-check", action="store_true", help="skip the post-run data-parallel correctness check")
    args = ap.parse_args(argv)
    if args.image <= 0:
        args.image = 299 if args.model == "inception" else 224       # BASELINE.json: 299x299 for Inception-v3
    return args


def metric_name(model: str) -> str:
    pretty = {"resnet34": "ResNet-34", "resnet18": "ResNet-18", "vgg16": "VGG-16", "vgg": "VGG-11-bn",
              "densenet": "DenseNet-121", "inception": "Inception-v3", "alexnet": "AlexNet",
              "squeezenet": "SqueezeNet-1.0"}.get(model, model)
    return "{} train img/s (max over ranks)".format(pretty)


def baseline_number(model: str, n: int):
    """BASELINE.md §5's measured baseline-(b) number for this model and GPU count (stock DDP, CUDA-graphed), or None.
    The reference publishes no throughput (BASELINE.json ``published`` is empty), so this is the only number
    ``vs_baseline`` can be taken against."""
    try:
        table = json.load(open(os.path.join(ROOT, "baseline", "ddp_baseline.json")))
        return float(table[model][str(n)])
    except Exception:  # noqa: BLE001 - no recorded baseline for this config
        return None


# ------------------------------------------------------------------------------------------ reference arm
def run_reference(args) -> None:
    """Unmodified reference code from baseline/_ref: models.initialize_model + mpi_tools.sync_params /
    mpi_avg_grads + the step order of main.py:146-156, CPU, torch.set_num_threads(1) (main.py:50-51).
    mpi4py itself is not installable offline; baseline/mpi4py_shim provides MPI.COMM_WORLD over gloo."""
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    rank = int(os.environ.get("RANK", "0"))
    if not os.path.exists(os.path.join(ref_dir, "mpi_tools.py")) and os.path.isdir("/root/reference") and rank == 0:
        import subprocess                                # baseline/_ref is git-ignored: re-create it where the source exists
        subprocess.run([sys.executable, os.path.join(ROOT, "baseline", "install_reference.py")], capture_output=True)
    if not os.path.exists(os.path.join(ref_dir, "mpi_tools.py")):
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": "baseline/_ref missing (run baseline/install_reference.py)"}))
        return
    import hashlib
    try:
        manifest = json.load(open(os.path.join(ROOT, "baseline", "REFERENCE_MANIFEST.json")))["files"]
        for f in ("mpi_tools.py", "models.py", "utils.py"):
            if hashlib.sha256(open(os.path.join(ref_dir, f), "rb").read()).hexdigest() != manifest[f]:
                raise RuntimeError("baseline/_ref/{} differs from the reference".format(f))
    except Exception as e:  # noqa: BLE001
        if rank == 0:
            print(json.dumps({"impl": "reference", "unavailable": "manifest check failed: {}".format(e)}))
        return
    sys.path[:0] = [os.path.join(ROOT, "baseline", "mpi4py_shim"), ref_dir]
    for m in ("utils", "models", "mpi_tools", "helpers"):
        sys.modules.pop(m, None)
    import warnings
    warnings.filterwarnings("ignore")
    import torch
    import torch.nn as nn
    import utils as rutils          # the reference's constants module: "edit utils.py" == set attributes
    import models as rmodels
    import mpi_tools as rtools
    from mpi4py import MPI
    comm = MPI.COMM_WORLD
    size = comm.Get_size()
    comm._ensure()
    torch.set_num_threads(1)                                   # main.py:51
    device = "cpu"                                             # main.py:50
    per_rank = int(os.environ.get("MPB_REF_BATCH", "8"))
    rutils.MODEL_NAME, rutils.NUM_CLASSES, rutils.BATCH_SIZE = args.model, args.classes, per_rank
    torch.manual_seed(0)
    model, _ = rmodels.initialize_model(rutils.MODEL_NAME, rutils.NUM_CLASSES, rutils.FEATURE_EXTRACT,
                                        use_pretrained=False)   # pretrained weights need the network
    optimizer = torch.optim.Adam(model.parameters(), lr=rutils.LR)
    rtools.sync_params(model)
    model.to(device)
    criterion = nn.CrossEntropyLoss()
    g = torch.Generator().manual_seed(1 + rank)
    images = torch.randn(per_rank, 3, args.image, args.image, generator=g)
    labels = torch.randint(0, args.classes, (per_rank,), generator=g)
    model.train()

    def step():
        logits = model(images.float())
        loss = criterion(logits, labels)
        optimizer.zero_grad()
        loss.backward()
        rtools.mpi_avg_grads(model)
        optimizer.step()
        return loss.detach().item()

    for _ in range(args.warmup):
        step()
    comm.Barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    comm.Barrier()
    dt = time.perf_counter() - t0
    secs = comm.reduce(dt, op=MPI.SUM, root=0)
    if rank == 0:
        secs_max = dt if size == 1 else max(dt, secs / size)
        value = per_rank * size * args.steps / secs_max
        in_bytes = images.numel() * 4 + labels.numel() * 8
        print(json.dumps({
            "impl": "reference", "metric": metric_name(args.model), "value": value, "unit": "img/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": secs_max / args.steps * 1e3,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32",
            "data": "synthetic", "gpu_launches": 0,
            # the reference loop IS end to end: host tensors in, loss.item() on the host every step (main.py:146-156);
            # its device is the CPU, so "H2D"/"D2H" are the host tensors it consumes / the scalar it produces per step
            "e2e": {"value": value, "unit": "img/s", "h2d_bytes_per_step": int(in_bytes), "d2h_bytes_per_step": 4,
                    "ms_per_step": secs_max / args.steps * 1e3,
                    "note": "same timed region as `value`: the reference has no device-resident fast path"},
            "config": {"model": args.model, "global_batch": per_rank * size, "seq_len": args.image,
                       "classes": args.classes, "parallelism": "dp{}".format(size),
                       "device": "cpu (reference hard-codes device='cpu', 1 torch thread)",
                       "note": "unmodified /root/reference sources from baseline/_ref via its own API (models.initialize_model, "
                               "mpi_tools.sync_params/mpi_avg_grads, main.py step order); mpi4py replaced by a gloo-backed shim "
                               "because no MPI exists offline; per-rank batch {} instead of 256 to bound wall time "
                               "(measured CPU img/s at batch 8 / 64 / 256: 10.1 / 9.7 / 8.7, BASELINE.md §5)".format(per_rank)}}))


# ------------------------------------------------------------------------------------------ baseline (b): stock DDP
def run_ddp(args) -> None:
    """Stock ``torchvision.models`` + ``torch.nn.parallel.DistributedDataParallel`` over NCCL + cuDNN, channels-last,
    bf16 autocast, ``torch.optim.Adam(fused=True)``: the step of /root/reference/main.py:146-156 as a competent
    PyTorch user would write it for this box.  None of this repository's models, kernels or engine is on this path.
    Timed twice: eager launches, and the whole step (forward, backward, NCCL all-reduce, Adam) replayed as ONE CUDA
    graph -- ``value`` is the faster of the two, both are reported."""
    # NCCL inside graph capture needs the watchdog's async error handling off; torchrun exports it as 1, so override
    os.environ["TORCH_NCCL_ASYNC_ERROR_HANDLING"] = "0"
    os.environ["NCCL_ASYNC_ERROR_HANDLING"] = "0"
    import torch
    import torch.distributed as dist
    import torchvision
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        if rank == 0:
            print(json.dumps({"impl": "ddp", "unavailable": "no CUDA device"}))
        return
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    torch.backends.cudnn.benchmark = True
    torch.manual_seed(0)
    ctor = {"resnet34": "resnet34", "resnet18": "resnet18", "vgg16": "vgg16", "vgg": "vgg11_bn", "alexnet": "alexnet",
            "densenet": "densenet121", "inception": "inception_v3", "squeezenet": "squeezenet1_0"}[args.model]
    kw = {"num_classes": args.classes}
    if args.model == "inception":
        kw.update(aux_logits=True, init_weights=False)
    net = getattr(torchvision.models, ctor)(**kw).to(dev).to(memory_format=torch.channels_last)
    B, K, W = args.batch, args.steps, args.warmup
    g = torch.Generator(device=dev).manual_seed(1 + rank)
    xs = [torch.randn(B, 3, args.image, args.image, device=dev, generator=g).contiguous(memory_format=torch.channels_last)
          for _ in range(4)]
    ys = [torch.randint(0, args.classes, (B,), device=dev, generator=g) for _ in range(4)]
    crit = torch.nn.CrossEntropyLoss()
    side = torch.cuda.Stream(dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):                  # DDP built and warmed on the stream the graph will be captured on
        model = torch.nn.parallel.DistributedDataParallel(net, device_ids=[local], gradient_as_bucket_view=True) \
            if world > 1 else net
        opt = torch.optim.Adam(model.parameters(), lr=4e-4, fused=True, capturable=True)
    sx, sy = xs[0].clone(), ys[0].clone()

    def step(x, y):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = model(x)
            if args.model == "inception":
                loss = crit(out[0].float(), y) + 0.4 * crit(out[1].float(), y)
            else:
                loss = crit(out.float(), y)
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()
        return loss.detach()

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def timed(fn, k):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sync_all()
        e0.record()
        for i in range(k):
            loss = fn(i)
        e1.record()
        e1.synchronize()
        ms = e0.elapsed_time(e1)
        sync_all()
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms, loss

    from mpi_pytorch_b200.utils.timing import ClockSampler
    with torch.cuda.stream(side):
        for i in range(max(W, 11)):                # DDP rebuilds its buckets during the first iterations
            step(xs[i % 4], ys[i % 4])
        sampler = ClockSampler(local).start() if rank == 0 else None
        ms_eager, loss = timed(lambda i: step(xs[i % 4], ys[i % 4]), K)
        t0 = time.perf_counter()
        step(xs[0], ys[0])
        host_ms = (time.perf_counter() - t0) * 1e3
        # end to end, eager: pinned-host batch -> H2D every step, loss -> host every step
        hx = [x.cpu().pin_memory() for x in xs[:3]]
        hy = [y.cpu().pin_memory() for y in ys[:3]]

        def e2e_step(i):
            x = hx[i % 3].to(dev, non_blocking=True)
            y = hy[i % 3].to(dev, non_blocking=True)
            lo = step(x, y)
            _ = float(lo)
            return lo
        for i in range(2):
            e2e_step(i)
        ms_e2e_eager, _ = timed(e2e_step, K)
    torch.cuda.current_stream(dev).wait_stream(side)
    res = {"ms_graph": None, "ms_e2e_graph": None, "graph_err": None, "loss": loss}

    def emit():
        clocks = sampler.stop() if sampler is not None else {}
        if rank != 0:
            return
        ms_graph, ms_e2e_graph = res["ms_graph"], res["ms_e2e_graph"]
        ms = min(m for m in (ms_eager, ms_graph) if m is not None)
        ms_e2e = min(m for m in (ms_e2e_eager, ms_e2e_graph) if m is not None)
        img = lambda m: None if m is None else B * world * K / (m / 1e3)      # noqa: E731
        in_bytes = hx[0].numel() * hx[0].element_size() + hy[0].numel() * hy[0].element_size()
        print(json.dumps({
            "impl": "ddp", "metric": metric_name(args.model), "value": img(ms), "unit": "img/s", "n_gpus": world,
            "steps": K, "warmup": max(W, 11), "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "gpu_launches": 0,
            "e2e": {"value": img(ms_e2e), "unit": "img/s", "h2d_bytes_per_step": int(in_bytes), "d2h_bytes_per_step": 4,
                    "ms_per_step": ms_e2e / K},
            "config": {"model": args.model, "global_batch": B * world, "seq_len": args.image, "classes": args.classes,
                       "parallelism": "dp{}".format(world), "optimizer": "adam(fused, capturable)",
                       "stack": "torchvision.models.{} + torch DDP (NCCL) + cuDNN(benchmark) + channels_last + bf16 autocast"
                                .format(ctor),
                       "eager_img_s": img(ms_eager), "graphed_img_s": img(ms_graph),
                       "eager_e2e_img_s": img(ms_e2e_eager), "graphed_e2e_img_s": img(ms_e2e_graph),
                       "graph_error": res["graph_err"], "graph_arm": res.get("graph_arm", "the eager step, captured"),
                       "host_ms_per_eager_step": host_ms,
                       "final_loss": float(res["loss"]) if res["graph_err"] is None else None,
                       "l2": "per-step working set >> 126 MB L2; 4 input batches rotate; no explicit flush"},
            "clocks": {k: clocks.get(k) for k in ("sm_mhz", "sm_max_mhz", "reasons")}}), flush=True)

    if os.environ.get("MPB_DDP_GRAPH", "1") != "0":
        import signal

        def on_alarm(_sig, _frm):          # a capture / replay that never returns: report the eager arm and leave
            res["graph_err"] = "graph arm did not finish within {} s".format(limit)
            emit()
            os._exit(0)
        limit = int(os.environ.get("MPB_DDP_GRAPH_TIMEOUT", "150"))
        signal.signal(signal.SIGALRM, on_alarm)
        signal.alarm(limit)
        try:
            graph = torch.cuda.CUDAGraph()
            gstep = step
            if world > 1:
                # DDP's reducer is not capturable in this torch build (whole-step capture through the DDP wrapper
                # never returned at N = 2).  The graphed arm therefore keeps model, optimizer, cuDNN kernels and NCCL
                # but exchanges the gradients the plain way: every p.grad is a view of ONE flat buffer and a single
                # ncclAllReduce(AVG) over it is captured right after backward (87 MB at ~700 GB/s is ~0.2 ms of a
                # 24 ms step, so the lost overlap is far below the launch overhead the graph removes).
                params = [p for p in net.parameters() if p.requires_grad]
                flat = torch.zeros(sum(p.numel() for p in params), device=dev)
                off = 0
                for p in params:
                    p.grad = flat[off:off + p.numel()].as_strided(p.shape, p.stride())
                    off += p.numel()

                def gstep(x, y):
                    with torch.autocast("cuda", dtype=torch.bfloat16):
                        out = net(x)
                        if args.model == "inception":
                            loss = crit(out[0].float(), y) + 0.4 * crit(out[1].float(), y)
                        else:
                            loss = crit(out.float(), y)
                    flat.zero_()
                    loss.backward()
                    dist.all_reduce(flat, op=dist.ReduceOp.AVG)
                    opt.step()
                    return loss.detach()
                res["graph_arm"] = "plain module + one flat NCCL all-reduce(AVG) + fused Adam, whole step in one CUDA graph"
            with torch.cuda.stream(side):
                for i in range(3):
                    gstep(xs[i % 4], ys[i % 4])
                sync_all()
                with torch.cuda.graph(graph, stream=side):
                    sloss = gstep(sx, sy)

                def replay(i):
                    sx.copy_(xs[i % 4], non_blocking=True)
                    sy.copy_(ys[i % 4], non_blocking=True)
                    graph.replay()
                    return sloss

                def replay_e2e(i):
                    sx.copy_(hx[i % 3], non_blocking=True)
                    sy.copy_(hy[i % 3], non_blocking=True)
                    graph.replay()
                    _ = float(sloss)
                    return sloss
                for i in range(3):
                    replay(i)
                res["ms_graph"], res["loss"] = timed(replay, K)
                for i in range(2):
                    replay_e2e(i)
                res["ms_e2e_graph"], _ = timed(replay_e2e, K)
        except Exception as e:  # noqa: BLE001 - report the eager numbers and why the graph arm is missing
            res["graph_err"] = repr(e)[:300]
            res["ms_graph"] = res["ms_e2e_graph"] = None
        signal.alarm(0)
    emit()
    if world > 1:
        # tearing the NCCL communicator down after it took part in a graph capture never returned here (N = 2):
        # the result is printed, every rank leaves without the collective destroy
        sys.stdout.flush()
        os._exit(0)
    if world > 1:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------------------ our arm
def dp_check(tr, batch) -> dict:
    """Driver-visible proof that the multi-GPU step is numerically the reference's (mpi_tools.py:30-37 then
    main.py:155): compute this rank's local gradients once, then from one saved state apply them twice -- (A) the
    product path, peer-memory kernels that fuse sum, 1/W and Adam; (B) ``dist.all_reduce`` over NCCL, divide by W,
    the torch Adam formula -- and compare the resulting fp32 parameters (expected: a few float ulps)."""
    import torch
    import torch.distributed as dist
    from mpi_pytorch_b200.ops import reference as ref
    eng, W = tr.engine, tr.world.size
    eng.gather_optimizer_state()
    n = eng.layout.total
    saved = {"p": eng.param_arena[:n].clone(), "step": eng.step_count,
             "st": {k: v.clone() for k, v in eng.state.items()}}

    def restore():
        eng.param_arena[:n].copy_(saved["p"])
        for k, v in eng.state.items():
            v.copy_(saved["st"][k])
        eng.set_step(saved["step"])
        eng.refresh_shadow()

    with eng.suspended():                                          # local gradients of this rank's batch, ONCE
        loss = tr.forward_loss(*batch)
        eng.grad_arena.zero_()
        loss.backward()
    g_local = eng.grad_arena[:n].clone()
    eng.reduce_and_step()                                          # (A) the product's kernels on those gradients
    torch.cuda.synchronize(tr.device)
    pa = eng.param_arena[:n].clone()
    shadow_ok = True
    if eng.shadow_arena is not None:
        shadow_ok = bool(torch.equal(eng.shadow_arena[:n], pa.to(eng.shadow_arena.dtype)))
    restore()
    g = g_local.clone()                                            # (B) NCCL + the torch formula on the SAME gradients
    if eng.wire_arena is not None:
        g = g.to(torch.bfloat16).float()                           # the bf16 wire rounds each rank's contribution once
    if W > 1:
        dist.all_reduce(g)
        g.div_(W)
    pb = saved["p"].clone()
    h = eng.hyper
    if h.kind == "adam":
        ref.adam_step(pb, g, saved["st"]["exp_avg"].clone()[:n], saved["st"]["exp_avg_sq"].clone()[:n],
                      saved["step"] + 1, h.lr, h.beta1, h.beta2, h.eps, h.weight_decay)
    else:
        ref.sgd_step(pb, g, saved["st"]["momentum_buffer"].clone()[:n], saved["step"] + 1, h.lr, h.momentum, h.weight_decay)
    diff = float((pa - pb).abs().max())
    upd = float((pa - saved["p"]).abs().max())
    digest = pa.view(torch.int32).to(torch.int64).sum().reshape(1)
    same = True
    if W > 1:
        all_d = [torch.zeros_like(digest) for _ in range(W)]
        dist.all_gather(all_d, digest)
        same = all(int(d) == int(all_d[0]) for d in all_d)
    restore()
    return {"param_digest_equal_across_ranks": bool(same), "vs_nccl_max_abs": diff, "update_max_abs": upd,
            "bf16_shadow_exact": bool(shadow_ok)}


def allreduce_bandwidth(tr, iters: int = 10) -> dict:
    """Second half of BASELINE.json's metric: bus bandwidth of the fused gradient all-reduce + Adam kernels on the
    bucket schedule the model actually uses, timed on the device in isolation (max over ranks).
    bus GB/s = 2 (W-1)/W x gradient bytes / t, NCCL's convention."""
    import torch
    from mpi_pytorch_b200.utils.timing import max_over_ranks
    eng, W = tr.engine, tr.world.size
    if W == 1 or not hasattr(eng.coll, "launches"):
        return {}
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    out = {"buckets_mb": [round(b.numel * eng.grad_arena.element_size() / 2 ** 20, 2) for b in eng.layout.buckets]}
    for name, fused in (("fused_adam", True), ("avg_only", False)):
        def once():
            for b in eng.layout.buckets:
                if fused:
                    eng._launch(b)
                else:
                    eng.coll.all_reduce_avg_(eng.layout.bucket_view(eng.grad_arena, b), b.algo)
            if fused:
                eng._in_flight = []
                torch.cuda.current_stream(tr.device).wait_stream(eng.comm_stream)
        for _ in range(3):
            once()
        torch.cuda.synchronize(tr.device)
        eng.coll.barrier()
        e0.record()
        for _ in range(iters):
            once()
        e1.record()
        e1.synchronize()
        ms = max_over_ranks(e0.elapsed_time(e1), tr.device) / iters
        nbytes = eng.layout.total * eng.grad_arena.element_size()
        out[name + "_ms"] = ms
        out[name + "_bus_gbs"] = 2 * (W - 1) / W * nbytes / (ms / 1e3) / 1e9
    out["grad_bytes"] = eng.layout.total * eng.grad_arena.element_size()
    out["frac_of_770_nvlink"] = out["fused_adam_bus_gbs"] / 770.0
    return out


def run_ours(args) -> None:
    import torch
    from mpi_pytorch_b200 import ops
    from mpi_pytorch_b200.data.synthetic import HostStagedShard
    from mpi_pytorch_b200.engine.trainer import Trainer
    from mpi_pytorch_b200.parallel import bootstrap
    from mpi_pytorch_b200.utils.config import Settings
    from mpi_pytorch_b200.utils.timing import ClockSampler, max_over_ranks

    world = bootstrap.init_world("auto")
    cuda = world.device.type == "cuda"
    batch, image, classes, model = args.batch, args.image, args.classes, args.model
    if not cuda:                                   # CPU smoke so `python bench.py` also works on the build box
        batch, image, classes = 4, 64, 10
    s = Settings(MODEL_NAME=model, NUM_CLASSES=classes, BATCH_SIZE=batch, HEIGHT=image, WIDTH=image,
                 USE_PRETRAINED=False, FEATURE_EXTRACT=False, DTYPE="bf16", LR=4e-4, OPTIMIZER="adam",
                 ALLREDUCE_ALGO=os.environ.get("MPB_ALLREDUCE_ALGO", "auto"),
                 GRAD_DTYPE=os.environ.get("MPB_GRAD_DTYPE", "fp32"), SEED=0)
    tr = Trainer(s, world)
    tr.sync_params()
    tr.model.train()
    shard = tr.make_shard(total=batch * world.size * 4, shuffle=False)
    dev_batches = list(iter(shard))                # 4 distinct device-resident batches, reused round-robin
    K, W = args.steps, args.warmup
    dev = world.device
    use_graph = cuda and tr.graph_capable() and os.environ.get("MPB_CUDA_GRAPH", "1") != "0"
    step_fn = tr.train_step
    if use_graph:
        try:
            tr.train_step_graphed(*dev_batches[0])
            step_fn = tr.train_step_graphed
        except Exception as e:  # noqa: BLE001 - report and fall back to eager launches
            print("CUDA graph capture failed, running eagerly: {!r}".format(e), file=sys.stderr)
            use_graph = False

    def sync_all():
        if cuda:
            torch.cuda.synchronize(dev)
        if world.size > 1:
            import torch.distributed as dist
            dist.barrier()
        if cuda:
            torch.cuda.synchronize(dev)

    from mpi_pytorch_b200.engine.trainer import count_launches

    def launches() -> int:
        return count_launches(tr.engine)

    # ---- kernel-only: inputs already on the device, no host read-back inside the timed region
    for i in range(W):
        step_fn(*dev_batches[i % len(dev_batches)])
    sync_all()
    sampler = ClockSampler(world.local_rank).start() if (cuda and world.rank == 0) else None
    l0 = launches()
    if cuda:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
    t0 = time.perf_counter()
    for i in range(K):
        loss = step_fn(*dev_batches[i % len(dev_batches)])
    host_enqueue_ms = (time.perf_counter() - t0) * 1e3 / max(K, 1)     # CPU time to enqueue one step (no sync)
    if cuda:
        e1.record()
        e1.synchronize()
        ms = e0.elapsed_time(e1)
    else:
        ms = (time.perf_counter() - t0) * 1e3
    sync_all()
    n_launch = (launches() - l0) if not use_graph else K * tr.graph_launches_per_step
    clocks = sampler.stop() if sampler is not None else {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
    ms = max_over_ranks(ms, dev if world.backend == "nccl" else "cpu")
    value = batch * world.size * K / (ms / 1e3)
    final_loss = float(loss.item())

    # ---- end to end: pinned host batch -> H2D every step, loss -> D2H every step (lagged one step)
    e2e = None
    if not args.no_e2e:
        staged = HostStagedShard(shard, n_host_batches=3, device=dev)
        copy_stream = torch.cuda.Stream(dev) if cuda else None
        host_loss = torch.zeros(2, dtype=torch.float32).pin_memory() if cuda else torch.zeros(2)

        def fetch(i):
            if copy_stream is None:
                return staged.fetch(i)
            with torch.cuda.stream(copy_stream):
                b = staged.fetch(i)
            return b

        def run(n, timed):
            nxt = fetch(0)
            ev = None
            for i in range(n):
                if copy_stream is not None:
                    torch.cuda.current_stream(dev).wait_stream(copy_stream)
                cur = nxt
                nxt = fetch(i + 1)                          # overlaps with this step's compute
                lo = step_fn(*cur)
                if cuda:
                    for t in cur:
                        t.record_stream(torch.cuda.current_stream(dev))
                    if ev is not None:
                        ev.synchronize()                    # read step i-1's loss on the host
                        _ = float(host_loss[(i - 1) % 2])
                    host_loss[i % 2].copy_(lo, non_blocking=True)
                    ev = torch.cuda.Event()
                    ev.record()
                else:
                    _ = float(lo)
            if ev is not None:
                ev.synchronize()
                _ = float(host_loss[(n - 1) % 2])

        run(max(2, W // 2), False)
        sync_all()
        if cuda:
            e0.record()
        t0 = time.perf_counter()
        run(K, True)
        if cuda:
            e1.record()
            e1.synchronize()
            ms2 = e0.elapsed_time(e1)
        else:
            ms2 = (time.perf_counter() - t0) * 1e3
        sync_all()
        ms2 = max_over_ranks(ms2, dev if world.backend == "nccl" else "cpu")
        e2e = {"value": batch * world.size * K / (ms2 / 1e3), "unit": "img/s",
               "h2d_bytes_per_step": int(staged.bytes_per_step), "d2h_bytes_per_step": 4,
               "ms_per_step": ms2 / K}
    check = bw = None
    if cuda and not args.no_check:
        try:
            check = dp_check(tr, dev_batches[1])
            bw = allreduce_bandwidth(tr)
        except Exception as e:  # noqa: BLE001 - the headline number stands; say why the check is missing
            check = {"error": repr(e)[:300]}
    if world.rank == 0:
        fb = dict(ops.native().FALLBACKS) if (cuda and ops.native_enabled()) else {}
        base = baseline_number(model, world.size)
        out = {
            "metric": metric_name(model), "value": value, "unit": "img/s", "n_gpus": world.size, "steps": K, "warmup": W,
            "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": (value / base) if base else None,
            "dtype": "bf16" if cuda else "fp32", "data": "synthetic",
            "config": {"model": model, "global_batch": batch * world.size, "seq_len": image, "classes": classes,
                       "parallelism": "dp{}".format(world.size), "optimizer": "adam",
                       "l2": "per-step working set (activations, ~GBs) >> 126 MB L2; 4 input batches rotate; no explicit flush",
                       "device": str(dev), "allreduce": getattr(tr.engine.coll, "name", "?"),
                       "grad_wire_dtype": str(tr.engine.grad_arena.dtype).replace("torch.", ""),
                       "native_ops": bool(cuda and ops.native_enabled()), "library_fallbacks": fb,
                       "scratch_pool": dict(__import__("mpi_pytorch_b200.ops.scratch", fromlist=["stats"]).stats),
                       "final_loss": final_loss, "host_enqueue_ms_per_step": host_enqueue_ms,
                       "cuda_graph": bool(use_graph),
                       "baseline_b_img_s": base, "baseline_b": "stock torchvision + DDP/NCCL + cuDNN, CUDA-graphed "
                                                               "(bench.py --impl ddp; BASELINE.md §5)"},
            "clocks": {k: clocks.get(k) for k in ("sm_mhz", "sm_max_mhz", "reasons")},
            "e2e": e2e, "gpu_launches": int(n_launch), "dp_check": check,
            "allreduce_bus_gbs": (bw or {}).get("fused_adam_bus_gbs"), "allreduce": bw,
        }
        print(json.dumps(out))
    tr.close()
    bootstrap.shutdown()


# ------------------------------------------------------------------------------------------ config #5: eval pipeline
def run_pipeline(args) -> None:
    """BASELINE.json config #5: train a ResNet-18 checkpoint data-parallel on all ranks, then evaluate it with the
    rank-staged pipeline (3 preprocessing ranks + N-3 predictor ranks, /root/reference/evaluation_pipeline.py:162-199)
    and check the summed accuracy against a single-process evaluation of the same checkpoint."""
    import tempfile

    import torch
    import torch.distributed as dist
    from mpi_pytorch_b200 import ops
    from mpi_pytorch_b200.engine.checkpoint import checkpoint_path, make_checkpoint, save_checkpoint
    from mpi_pytorch_b200.engine.trainer import Trainer
    from mpi_pytorch_b200.models import initialize_model
    from mpi_pytorch_b200.parallel import bootstrap
    from mpi_pytorch_b200.parallel.pipeline import EvalPipeline
    from mpi_pytorch_b200.utils.config import Settings

    world = bootstrap.init_world("auto")
    dev, cuda = world.device, world.device.type == "cuda"
    n = int(os.environ.get("MPB_PIPE_IMAGES", "10000" if cuda else "96"))     # the reference's test_sample.csv has 10,000 rows
    raw_hw = (250, 170) if cuda else (40, 28)
    out_hw = (128, 128) if cuda else (32, 32)                                  # utils.py:33-34
    classes = int(os.environ.get("MPB_PIPE_CLASSES", "100" if cuda else "4"))
    train_steps = int(os.environ.get("MPB_PIPE_TRAIN_STEPS", str(args.steps * 15 if cuda else 30)))
    mb = int(os.environ.get("MPB_PIPE_MICROBATCH", "64" if cuda else "8"))
    B = 128 if cuda else 16
    dtype = torch.bfloat16 if cuda else torch.float32

    def make_frames(i0: int, i1: int, device) -> torch.Tensor:
        """Learnable stand-in for Herbarium JPEGs: noise plus a class-dependent colour cast (uint8 HWC)."""
        lab = labels[i0:i1].to(device)
        g = torch.Generator(device=device).manual_seed(7919 + i0)
        x = torch.randint(0, 96, (i1 - i0, raw_hw[0], raw_hw[1], 3), generator=g, dtype=torch.int32, device=device)
        for c in range(3):
            x[:, :, :, c] += ((lab % 3 == c).int() * (40 + 8 * (lab % 16).int())).view(-1, 1, 1)
            x[:, : raw_hw[0] // 2, :, c] += (((lab // 3) % 3 == c).int() * (6 * (lab % 24).int())).view(-1, 1, 1)
        return x.clamp_(0, 255).to(torch.uint8)

    from mpi_pytorch_b200.data.synthetic import synth_labels
    labels = synth_labels(0, torch.arange(n), classes)
    # The whole "decoded" test set is generated once, up front, on every rank's device (same seeds -> same frames);
    # the reader rank additionally keeps it in pinned host memory, which is what its `source` reads from, so every
    # message of the timed pipeline still starts with a host -> device copy of raw frames.
    CH = 500
    dev_frames = torch.cat([make_frames(i, min(i + CH, n), dev) for i in range(0, n, CH)])
    host_frames = None
    if world.rank == 0:
        host_frames = dev_frames.cpu()
        if cuda:
            host_frames = host_frames.pin_memory()

    def frames(i0: int, i1: int) -> torch.Tensor:
        return host_frames[i0:i1]
    s = Settings(MODEL_NAME="resnet18", NUM_CLASSES=classes, BATCH_SIZE=B, HEIGHT=out_hw[0], WIDTH=out_hw[1],
                 USE_PRETRAINED=False, DTYPE="bf16", LR=1e-3, SEED=0, VALIDATE=False)
    tr = Trainer(s, world)
    tr.sync_params()
    tr.model.train()
    # ---- phase 1: data-parallel training on the (train == eval) sample, preprocessed like the pipeline does
    g = torch.Generator().manual_seed(1234 + world.rank)
    t_train = time.perf_counter()
    last = None
    for step in range(train_steps):
        i0 = int(torch.randint(0, n - B + 1, (1,), generator=g))         # a window of B consecutive samples
        raw = dev_frames[i0:i0 + B]
        x = ops.normalize_u8(ops.resize_u8(raw, out_hw[0], out_hw[1], "bicubic"), dtype)
        if cuda:
            x = x.contiguous(memory_format=torch.channels_last)
        last = tr.train_step(x, labels[i0:i0 + B].to(dev))
    train_loss = float(last) if last is not None else float("nan")
    t_train = time.perf_counter() - t_train
    ckdir = os.environ.get("MPB_PIPE_CKPT_DIR", os.path.join(tempfile.gettempdir(), "mpb_pipe_ckpt"))
    ck = make_checkpoint(0, tr.model, tr.optimizer, train_loss)          # collective
    if world.rank == 0:
        save_checkpoint(ck, 0, "resnet18", ckdir, None, False)
    if world.size > 1:
        dist.barrier()
    tr.close()

    # ---- phase 2: the staged pipeline over all ranks
    def model_fn():
        m, _ = initialize_model("resnet18", classes, feature_extract=False, use_pretrained=False)
        m.load_state_dict(torch.load(checkpoint_path(ckdir, "resnet18"), map_location="cpu", weights_only=True)["state_dict"])
        if cuda:
            m = ops.prepare_inference(m.to(dev).to(memory_format=torch.channels_last), torch.bfloat16)
        return m
    coll = tr.engine.coll if (cuda and world.size > 1) else None      # the trainer's peer-memory runtime, reused
    pipe = EvalPipeline(world, n, raw_hw, out_hw, frames, labels, model_fn, micro_batch=mb, seed=0, dtype=dtype,
                        collectives=coll)
    acc = pipe.run()
    stats = pipe.stats
    busy = torch.tensor([stats["stage_busy_frac"]], dtype=torch.float64, device=dev)
    busy_all = [torch.zeros_like(busy) for _ in range(world.size)]
    if world.size > 1:
        dist.all_gather(busy_all, busy)
    else:
        busy_all = [busy]
    # ---- phase 3: single-process oracle on one rank (same preprocessing functions, same micro-batches)
    single = None
    if world.rank == world.size - 1:
        m = model_fn().to(dev).eval()
        correct = torch.zeros((), dtype=torch.int64, device=dev)
        with torch.no_grad():
            for i0 in range(0, n, mb):
                i1 = min(i0 + mb, n)
                x = ops.normalize_u8(ops.resize_u8(dev_frames[i0:i1], out_hw[0], out_hw[1], "bicubic"), dtype)
                if cuda:
                    x = x.contiguous(memory_format=torch.channels_last)
                correct += ops.count_correct(m(x), labels[i0:i1].to(dev))
        single = float(correct.item()) / n
    box = [single]
    if world.size > 1:
        dist.broadcast_object_list(box, src=world.size - 1)
    single = box[0]
    if world.rank == 0:
        roles = pipe.roles
        print(json.dumps({
            "task": "pipeline", "metric": "pipelined evaluation img/s (ResNet-18 checkpoint)", "value": stats["images_per_s"],
            "unit": "img/s", "n_gpus": world.size, "higher_is_better": True, "dtype": "bf16" if cuda else "fp32",
            "data": "synthetic (learnable colour-cast frames, train == eval sample as in main.py:108)",
            "accuracy": acc, "single_process_accuracy": single,
            "accuracy_matches_single_process": bool(abs(acc - single) <= max(2.0 / n, 0.002)),
            "elapsed_ms": stats["elapsed_ms"],
            "stage_busy_frac": {"reader": float(busy_all[roles.reader]), "resizer": float(busy_all[roles.resizer]),
                                "normalizer": float(busy_all[roles.normalizer]),
                                "predictors": [float(busy_all[p]) for p in roles.predictors]},
            "config": {"images": n, "raw_hw": raw_hw, "out_hw": out_hw, "micro_batch": mb, "classes": classes,
                       "roles": {"reader": roles.reader, "resizer": roles.resizer, "normalizer": roles.normalizer,
                                 "predictors": roles.predictors},
                       "train_steps": train_steps, "train_loss": train_loss, "train_s": t_train,
                       "p2p": "one 2-rank NCCL communicator per edge, pre-posted irecv double buffers"}}))
        assert abs(acc - single) <= max(2.0 / n, 0.002), (acc, single)
    bootstrap.shutdown()


def main(argv=None):
    args = parse_args(argv)
    if args.task == "pipeline":
        run_pipeline(args)
    elif args.impl == "reference":
        run_reference(args)
    elif args.impl == "ddp":
        run_ddp(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
