"""Device-timed BatchNorm(+ReLU/+residual) forward and backward passes at ResNet-34 activation sizes, with achieved
HBM bandwidth (bytes actually moved / time) against MEASURED_PEAKS.json's copy rate."""
import json
import os
import sys

import torch
import torch.nn as nn

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


def main():
    from mpi_pytorch_b200 import ops
    nat = ops.native()
    peak = 6463.0
    try:
        peak = float(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_copy_gbps"])
    except Exception:
        pass
    rows = []
    for (name, n, c, hw) in (("stem", 256, 64, 112), ("l1", 256, 64, 56), ("l2", 256, 128, 28), ("l3", 256, 256, 14),
                             ("l4", 256, 512, 7)):
        for res in (False, True):
            x = torch.randn(n, c, hw, hw, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
            x.requires_grad_(True)
            r = torch.randn_like(x).requires_grad_(True) if res else None
            bn = nn.BatchNorm2d(c).cuda()
            g = torch.randn_like(x)
            mb = x.numel() * 2 / 1e6
            out = nat.batch_norm_act(x, bn, r, True)
            t_f = timed(lambda: nat.batch_norm_act(x, bn, r, True))          # stats (x) + apply (x [+res] -> out)
            t_fb = timed(lambda: nat.batch_norm_act(x, bn, r, True).backward(g))
            f_bytes = mb * (1 + 2 + (1 if res else 0))
            b_bytes = mb * ((3 + 3 + 2) if res else (2 + 2 + 1))
            rows.append({"layer": name, "residual": res, "MB": round(mb, 1), "fwd_ms": round(t_f, 4),
                         "fwd_GBps": round(f_bytes / t_f, 0), "bwd_ms": round(t_fb - t_f, 4),
                         "bwd_GBps": round(b_bytes / max(t_fb - t_f, 1e-6), 0), "peak_GBps": peak})
            print(json.dumps(rows[-1]), flush=True)
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(rows, open("gpurun_out/bn_bench.json", "w"), indent=1)


if __name__ == "__main__":
    main()
