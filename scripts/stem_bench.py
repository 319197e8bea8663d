"""Device-timed 7x7/2 stem (space-to-depth form): s2d copy, fprop, wgrad at batch 256, 224^2."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def timed(fn, iters=10):
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
    from mpi_pytorch_b200.ops import tc
    n = int(os.environ.get("N", "256"))
    x = torch.randn(n, 3, 224, 224, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    w = (torch.randn(64, 3, 7, 7, device="cuda") * 0.1).bfloat16().contiguous(memory_format=torch.channels_last)
    y, stats, xs = tc.conv_fprop_stem(x, w, None, (3, 3), True)
    gy = torch.randn_like(y)
    print("fprop (s2d copy + filter gather + conv + stats) ms:", round(timed(lambda: tc.conv_fprop_stem(x, w, None, (3, 3), True)), 4))
    print("fprop without stats ms:", round(timed(lambda: tc.conv_fprop_stem(x, w, None, (3, 3), False)), 4))
    print("wgrad ms:", round(timed(lambda: tc.conv_wgrad_stem(gy, xs, x.shape, w.shape, (3, 3))), 4))
    flops = 2.0 * n * 112 * 112 * 64 * 147
    print("useful GFLOP per pass:", round(flops / 1e9, 1))


if __name__ == "__main__":
    main()
