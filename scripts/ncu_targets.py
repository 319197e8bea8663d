"""One launch (after a warm-up launch) of each kernel that had no ncu capture in round 1: max-pool fwd/bwd, softmax
cross-entropy, resize+normalise, flat Adam, the fused all-reduce+Adam kernel (W = 1: its local data path), the channel
concat copy and the strided-dgrad tensor-core launch.  Meant to be run under
  ncu --set full --clock-control none --import-source on -k regex:<pattern> ...
and, without ncu, prints CUDA-event timings of the same launches with achieved GB/s against the measured 6,463 GB/s."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_pytorch_b200 import ops  # noqa: E402
from mpi_pytorch_b200.ops import tc  # noqa: E402
from mpi_pytorch_b200.parallel import bootstrap  # noqa: E402
from mpi_pytorch_b200.parallel.collectives import Hyper  # noqa: E402
from mpi_pytorch_b200.parallel.peer import PeerCollectives  # noqa: E402

HBM = 6463.0


def timed(fn, reps=10):
    fn(); fn()
    torch.cuda.synchronize()
    if os.environ.get("MPB_NCU") == "1":               # under ncu: the two launches above are all that is wanted
        return 1.0
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
    best = 1e9
    for _ in range(reps):
        flush.zero_()                                   # > 126 MB: L2 holds nothing of the operands
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); e1.synchronize()
        best = min(best, e0.elapsed_time(e1))
    return best


def main():
    nat = ops.native()
    dev = torch.device("cuda")
    cl = torch.channels_last
    rows = []

    def add(name, ms, nbytes):
        rows.append({"kernel": name, "ms": ms, "GBs": nbytes / ms / 1e6, "frac_of_hbm_copy": nbytes / ms / 1e6 / HBM})

    x = torch.randn(256, 64, 112, 112, device=dev).to(torch.bfloat16).contiguous(memory_format=cl).requires_grad_(True)
    y = nat.max_pool2d(x, 3, 2, 1)
    gy = torch.randn_like(y)
    add("maxpool3x3_fwd (256x64x112^2 bf16)", timed(lambda: nat.max_pool2d(x, 3, 2, 1)), x.numel() * 2 + y.numel() * 3)
    add("maxpool3x3s2_bwd", timed(lambda: torch.autograd.grad(y, x, gy, retain_graph=True)), y.numel() * 3 + x.numel() * 2)
    for C in (1000, 64500):
        lg = torch.randn(256, C, device=dev).to(torch.bfloat16).requires_grad_(True)
        tgt = torch.randint(0, C, (256,), device=dev)
        add("softmax_ce fwd+dlogits (256x{})".format(C), timed(lambda: nat.cross_entropy(lg, tgt)), lg.numel() * 4)
    img = torch.randint(0, 255, (64, 250, 170, 3), device=dev, dtype=torch.uint8)
    add("resize bicubic u8 (64x250x170 -> 128^2)", timed(lambda: nat.resize_u8(img, 128, 128, "bicubic")), img.numel() + 64 * 128 * 128 * 3)
    add("resize+normalise bilinear -> bf16 NHWC", timed(lambda: nat.resize_normalize(img, 128, 128, "bilinear", torch.bfloat16, True)),
        img.numel() + 64 * 128 * 128 * 3 * 2)
    n = 21_797_672 // 64 * 64
    p, g, m, v = (torch.randn(n, device=dev) for _ in range(4))
    v.abs_()
    sh = torch.empty(n, dtype=torch.bfloat16, device=dev)
    add("adam_flat (21.8 M params, fp32 + bf16 shadow)", timed(lambda: nat.adam_flat(p, g, m, v, sh, 2, 4e-4, 0.9, 0.999, 1e-8, 0.0)),
        n * (4 * 4 + 3 * 4 + 2))
    pc = PeerCollectives(bootstrap.init_world("cuda"))
    n2 = 8 << 20
    g2, p2, s2 = pc.alloc(n2, torch.float32), pc.alloc(n2, torch.float32), pc.alloc(n2, torch.bfloat16)
    st = {"exp_avg": torch.zeros(n2, device=dev), "exp_avg_sq": torch.zeros(n2, device=dev)}
    g2.normal_()
    add("allreduce_kernel<1> fused Adam, 32 MB bucket (local data path)",
        timed(lambda: pc.all_reduce_update_(g2, p2, st, s2, Hyper("adam", lr=1e-4), 2, "twoshot")), n2 * (4 * 4 + 3 * 4 + 2))
    a, b = (torch.randn(64, c, 28, 28, device=dev).to(torch.bfloat16).contiguous(memory_format=cl) for c in (256, 32))
    add("channel concat copy (256+32 ch @28^2 x64)", timed(lambda: nat.cat_channels([a, b])), (a.numel() + b.numel()) * 4)
    dy = torch.randn(256, 128, 28, 28, device=dev).to(torch.bfloat16).contiguous(memory_format=cl)
    w = torch.randn(128, 64, 3, 3, device=dev).to(torch.bfloat16).contiguous(memory_format=cl)
    ms = timed(lambda: tc.conv_dgrad(dy, w, (256, 64, 56, 56), (2, 2), (1, 1)))
    rows.append({"kernel": "tc_gemm_kernel strided dgrad 3x3/2 128->64 @28^2->56^2, one launch", "ms": ms,
                 "TFLOPs": 2 * 256 * 28 * 28 * 128 * 64 * 9 / ms / 1e9})
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rows, open("gpurun_out/ncu_targets_timing.json", "w"), indent=1)
    for r in rows:
        print(json.dumps(r))


if __name__ == "__main__":
    main()
