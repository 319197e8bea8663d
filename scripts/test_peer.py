"""Multi-GPU check of the peer-memory collectives against torch.distributed (NCCL).
Run:  torchrun --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 scripts/test_peer.py
Also measures all-reduce bus bandwidth (device-timed, max over ranks) for both backends."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_pytorch_b200.ops import reference as ref  # noqa: E402
from mpi_pytorch_b200.parallel import bootstrap  # noqa: E402
from mpi_pytorch_b200.parallel.collectives import Hyper  # noqa: E402
from mpi_pytorch_b200.parallel.peer import PeerCollectives  # noqa: E402


def main():
    w = bootstrap.init_world("cuda")
    pc = PeerCollectives(w)
    W, r = w.size, w.rank
    ok = True
    results = {"world": W, "checks": [], "bandwidth": []}
    results["nvls"] = bool(pc.nvls)
    for n, algo in ((1024, "oneshot"), (64 * 1024, "oneshot"), (1 << 20, "twoshot"), (8 << 20, "twoshot"), (1 << 16, "auto"),
                    (1 << 20, "nvls"), (8 << 20, "nvls")):
        n = n // (4 * W) * (4 * W)
        g = pc.alloc(n, torch.float32)
        torch.manual_seed(100 + r)
        g.normal_()
        want = g.clone()
        dist.all_reduce(want)
        want /= W
        pc.all_reduce_avg_(g, algo)
        torch.cuda.synchronize()
        err = (g - want).abs().max().item()
        ok &= err < 1e-5
        results["checks"].append({"op": "avg", "n": n, "algo": algo, "max_err": err})
        # fused adam: every rank must end with identical params == adam(avg grad)
        p = pc.alloc(n, torch.float32); sh = pc.alloc(n, torch.bfloat16)
        torch.manual_seed(7); p.normal_()
        g.normal_()
        gavg = g.clone(); dist.all_reduce(gavg); gavg /= W
        m = torch.zeros(n, device=w.device); v = torch.zeros(n, device=w.device)
        p2 = p.clone(); m2 = torch.zeros_like(p2); v2 = torch.zeros_like(p2)
        pc.all_reduce_update_(g, p, {"exp_avg": m, "exp_avg_sq": v}, sh, Hyper("adam", lr=1e-3), 1, algo)
        torch.cuda.synchronize(); dist.barrier()
        ref.adam_step(p2, gavg, m2, v2, 1, 1e-3)
        err = (p - p2).abs().max().item()
        e2 = (sh.float() - p.bfloat16().float()).abs().max().item()
        ok &= err < 1e-5 and e2 == 0
        results["checks"].append({"op": "adam", "n": n, "algo": algo, "max_err": err, "shadow_err": e2})
    # bf16 wire format: fp32 accumulation in rank order, one rounding
    for n in (4096, 1 << 20, 8 << 20):
        n = n // (8 * W) * (8 * W)
        gb = pc.alloc(n, torch.bfloat16)
        torch.manual_seed(200 + r)
        gb.copy_(torch.randn(n, device=w.device))
        want = gb.float()
        dist.all_reduce(want)
        want = (want / W).bfloat16()
        pc.all_reduce_avg_(gb)
        torch.cuda.synchronize(); dist.barrier()
        err = (gb.float() - want.float()).abs().max().item()
        ok &= err <= 2e-2                                  # at most one bf16 ulp of an O(1) value (summation order)
        results["checks"].append({"op": "avg_bf16", "n": n, "max_err": err})
    # broadcast
    b = pc.alloc(1 << 18, torch.float32)
    b.fill_(float(r + 1))
    pc.broadcast_(b, 0)
    torch.cuda.synchronize()
    ok &= bool((b == 1.0).all())
    c = torch.tensor(r + 5, dtype=torch.int64, device=w.device)
    pc.all_reduce_sum_i64(c)
    torch.cuda.synchronize()
    ok &= int(c) == sum(range(5, 5 + W))
    # bandwidth: algorithmic bus bandwidth = 2*(W-1)/W * bytes / t
    for mb in (0.25, 1, 8, 32, 128, 512):
        n = int(mb * 2 ** 20 / 4) // (4 * W) * (4 * W)
        g = pc.alloc(n, torch.float32)
        t = torch.empty(n, device=w.device)
        for name, fn in (("peer_twoshot", lambda: pc.all_reduce_avg_(g, "twoshot" if mb > 0.25 else "oneshot")),
                         ("peer_auto(nvls)" if pc.nvls else "peer_auto", lambda: pc.all_reduce_avg_(g, "auto")),
                         ("nccl", lambda: dist.all_reduce(t))):
            for _ in range(5):
                fn()
            torch.cuda.synchronize(); dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            iters = 20
            e0.record()
            for _ in range(iters):
                fn()
            e1.record(); e1.synchronize()
            ms = torch.tensor([e0.elapsed_time(e1) / iters], device=w.device)
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            bus = 2 * (W - 1) / W * n * 4 / (ms.item() * 1e-3) / 1e9
            results["bandwidth"].append({"impl": name, "MB": mb, "us": ms.item() * 1e3, "bus_GBs": bus})
    for mb in (8, 32, 128):                                # bf16 all-reduce: same element counts as the fp32 rows
        n = int(mb * 2 ** 20 / 4) // (8 * W) * (8 * W)
        gb = pc.alloc(n, torch.bfloat16)
        tb = torch.empty(n, device=w.device, dtype=torch.bfloat16)
        for name, fn in (("peer_bf16", lambda: pc.all_reduce_avg_(gb)), ("nccl_bf16", lambda: dist.all_reduce(tb))):
            for _ in range(5):
                fn()
            torch.cuda.synchronize(); dist.barrier()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                fn()
            e1.record(); e1.synchronize()
            ms = torch.tensor([e0.elapsed_time(e1) / 20], device=w.device)
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            results["bandwidth"].append({"impl": name, "MB_fp32_equiv": mb, "us": ms.item() * 1e3,
                                         "bus_GBs": 2 * (W - 1) / W * n * 2 / (ms.item() * 1e-3) / 1e9})
    flag = torch.tensor([1 if ok else 0], device=w.device)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    if r == 0:
        results["ok"] = bool(flag.item())
        print(json.dumps(results))
        os.makedirs("gpurun_out", exist_ok=True)
        with open("gpurun_out/peer_{}gpu.json".format(W), "w") as fh:
            json.dump(results, fh, indent=1)
    bootstrap.shutdown()
    sys.exit(0 if flag.item() else 1)


if __name__ == "__main__":
    main()
