"""Run fprop, dgrad and wgrad of ONE conv layer a few times (the target of `ncu -k regex:... -c N` captures).
  python scripts/one_layer.py 256 64 56 56 64 3 1 1     # N C H W K R stride pad
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    n, c, h, w, k, r, stride, pad = (int(v) for v in (sys.argv[1:9] if len(sys.argv) >= 9 else "256 64 56 56 64 3 1 1".split()))
    iters = int(os.environ.get("ITERS", "3"))
    from mpi_pytorch_b200.ops import tc
    torch.manual_seed(0)
    x = torch.randn(n, c, h, w, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    wt = (torch.randn(k, c, r, r, device="cuda") * 0.05).bfloat16().contiguous(memory_format=torch.channels_last)
    st, pd = (stride, stride), (pad, pad)
    for _ in range(iters):
        y, _ = tc.conv_fprop(x, wt, None, st, pd, True)
        gy = torch.randn_like(y)
        tc.conv_dgrad(gy, wt, x.shape, st, pd)
        tc.conv_wgrad(gy, x, wt.shape, st, pd)
    torch.cuda.synchronize()
    print("ok", tuple(y.shape))


if __name__ == "__main__":
    main()
