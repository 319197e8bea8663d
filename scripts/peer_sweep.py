"""Measure the fused all-reduce kernels over (bucket bytes) x (algorithm) x (grid) at the launched world size and
write the per-(world, bytes) choice table the engine loads (mpi_pytorch_b200/parallel/algo_table.json).

  torchrun --nproc-per-node W --master-addr 127.0.0.1 --master-port 29544 scripts/peer_sweep.py [--quick]

Every number is device-timed (CUDA events around CUDA-graph replays of 10 back-to-back launches, so the host launch
path is not in it), max over ranks.  Two variants per point: `avg` (gradients averaged in place) and `adam` (the
product path: reduce-scatter -> 1/W -> Adam on the owner slice -> all-gather of fp32 params + bf16 shadow).
bus GB/s = 2 (W-1)/W x fp32 gradient bytes / t (NCCL's convention; roofline 770 GB/s per direction per GPU).
Replaces the fixed 256 KB crossover (VERDICT r1 #3b); reference site being replaced: mpi_tools.py:30-37.
"""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_pytorch_b200.parallel import bootstrap  # noqa: E402
from mpi_pytorch_b200.parallel.collectives import Hyper  # noqa: E402
from mpi_pytorch_b200.parallel.peer import PeerCollectives  # noqa: E402

INNER = 10


def timed_graph(fn, dev, reps=3):
    """ms per call of `fn`, measured on replays of a graph holding INNER calls; max over ranks."""
    side = torch.cuda.Stream(dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):
        for _ in range(2):
            fn()
    torch.cuda.current_stream(dev).wait_stream(side)
    torch.cuda.synchronize(dev)
    dist.barrier()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(INNER):
            fn()
    g.replay()
    torch.cuda.synchronize(dev)
    dist.barrier()
    best = 1e30
    for _ in range(reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        e1.synchronize()
        best = min(best, e0.elapsed_time(e1) / INNER)
        dist.barrier()
    t = torch.tensor([best], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    quick = "--quick" in sys.argv
    w = bootstrap.init_world("cuda")
    pc = PeerCollectives(w)
    W, dev = w.size, w.device
    sizes_mb = [0.0625, 0.25, 1, 2, 4, 8, 16, 32, 64, 128, 512] if not quick else [0.25, 2, 32]
    grids = [16, 32, 64, 128] if not quick else [32, 64]
    rows = []
    hyper = Hyper("adam", lr=1e-4)
    max_n = int(max(sizes_mb) * 2 ** 20 / 4) // (8 * W) * (8 * W)
    g_all, p_all = pc.alloc(max_n, torch.float32), pc.alloc(max_n, torch.float32)
    sh_all, wire_all = pc.alloc(max_n, torch.bfloat16), pc.alloc(max_n, torch.bfloat16)
    m_all, v_all = torch.zeros(max_n, device=dev), torch.zeros(max_n, device=dev)
    t_all = torch.empty(max_n, device=dev)
    g_all.normal_()
    pc.step_dev = torch.ones(1, dtype=torch.int32, device=dev)
    for mb in sizes_mb:
        n = int(mb * 2 ** 20 / 4) // (8 * W) * (8 * W)
        g, p, sh, wire = g_all[:n], p_all[:n], sh_all[:n], wire_all[:n]
        st = {"exp_avg": m_all[:n], "exp_avg_sq": v_all[:n]}
        nbytes = n * 4
        bus = lambda ms: 2 * (W - 1) / W * nbytes / (ms * 1e-3) / 1e9      # noqa: E731
        algos = ["twoshot"] + (["nvls"] if pc.nvls else [])
        for grid in grids:
            if n <= pc.lib.mpb_oneshot_max_elems(grid):
                algos_g = ["oneshot"] + algos
            else:
                algos_g = algos
            for algo in algos_g:
                spec = "{}:{}".format(algo, grid)
                ms_avg = timed_graph(lambda: pc.all_reduce_avg_(g, spec), dev)
                ms_adam = timed_graph(lambda: pc.all_reduce_update_(g, p, st, sh, hyper, 1, spec), dev)
                row = {"MB": mb, "algo": algo, "grid": grid, "avg_us": ms_avg * 1e3, "avg_bus_GBs": bus(ms_avg),
                       "adam_us": ms_adam * 1e3, "adam_bus_GBs": bus(ms_adam)}
                if algo != "oneshot":
                    ms_w = timed_graph(lambda: pc.all_reduce_update_(g, p, st, sh, hyper, 1, spec, wire=wire), dev)
                    row.update(adam_bf16wire_us=ms_w * 1e3, adam_bf16wire_bus_GBs=bus(ms_w))
                rows.append(row)
        ms_nccl = timed_graph(lambda: dist.all_reduce(t_all[:n]), dev)
        rows.append({"MB": mb, "algo": "nccl", "grid": 0, "avg_us": ms_nccl * 1e3, "avg_bus_GBs": bus(ms_nccl)})
    # choice table: per size the fastest fused-Adam (algo, grid); ties within 3 % go to the smaller grid, which
    # leaves more SMs to the backward kernels the collective overlaps with
    table = []
    for mb in sizes_mb:
        # grids above 64 CTAs are measured but not chosen: a comm CTA (512 threads, ~116 registers) owns its SM, and
        # the collective shares the GPU with the backward kernels it overlaps
        cand = [r for r in rows if r["MB"] == mb and r["algo"] != "nccl" and r["grid"] <= 64]
        best = min(c["adam_us"] for c in cand)
        pick = min((c for c in cand if c["adam_us"] <= best * 1.03), key=lambda c: c["grid"])
        table.append([int(mb * 2 ** 20), pick["algo"], pick["grid"]])
    if w.rank == 0:
        out = {"world": W, "nvls": bool(pc.nvls), "inner_launches_per_graph": INNER, "rows": rows,
               "table": {str(W): table}}
        os.makedirs("gpurun_out", exist_ok=True)
        with open("gpurun_out/peer_sweep_{}gpu.json".format(W), "w") as fh:
            json.dump(out, fh, indent=1)
        for r in rows:
            print(json.dumps(r))
        print("TABLE", json.dumps({str(W): table}))
    bootstrap.shutdown()


if __name__ == "__main__":
    main()
