"""Per-case diagnostic for the tcgen05 GEMM / implicit-GEMM conv kernels.

`python scripts/tc_diag.py` runs every case in its own subprocess with a timeout (a wedged mbarrier
pipeline must not take the whole GPU call down) and prints one JSON line per case with the max
abs / relative error against an fp32 torch oracle computed from the same bf16-rounded inputs, plus a
device-timed TFLOP/s figure.  `--case NAME` runs a single case in-process.
"""
import argparse
import json
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

GEMM_CASES = {
    # name: (M, N, K, a_mn, b_mn)
    "gemm_tn_small": (128, 64, 64, 0, 0),
    "gemm_tn_k256": (128, 128, 256, 0, 0),
    "gemm_tn_head": (256, 1000, 512, 0, 0),
    "gemm_tn_big": (4096, 4096, 4096, 0, 0),
    "big_a_mn": (4096, 4096, 4096, 1, 0),      # operand-majorness sweep at a compute-bound size
    "big_b_mn": (4096, 4096, 4096, 0, 1),
    "big_ab_mn": (4096, 4096, 4096, 1, 1),
    "gemm_bmn": (256, 512, 1000, 0, 1),        # linear dX shape
    "gemm_amn_bmn": (1000, 512, 256, 1, 1),    # linear dW shape
    "gemm_amn_bmn_small": (128, 64, 64, 1, 1),
    # GEMM twins of the conv layers (same M, N, K, tiled TMA instead of im2col TMA)
    "twin_l1": (802816, 64, 576, 0, 0),
    "twin_l2": (200704, 128, 1152, 0, 0),
    "twin_l3": (50176, 256, 2304, 0, 0),
    "twin_l4": (12544, 512, 4608, 0, 0),
}
CONV_CASES = {
    # name: (N, C, H, W, K, R, stride, pad)
    "conv_1x1": (2, 64, 8, 8, 64, 1, 1, 0),
    "conv_3x3_small": (2, 64, 8, 8, 64, 3, 1, 1),
    "conv_l1": (8, 64, 56, 56, 64, 3, 1, 1),
    "conv_l2_s2": (8, 64, 56, 56, 128, 3, 2, 1),
    "conv_l2_ds": (8, 64, 56, 56, 128, 1, 2, 0),
    "conv_l3": (8, 256, 14, 14, 256, 3, 1, 1),
    "conv_l4": (16, 512, 7, 7, 512, 3, 1, 1),
    "conv_nopad": (4, 64, 17, 17, 128, 3, 1, 0),
    # full-size ResNet-34 layers at batch 256 (performance cases)
    "full_l1": (256, 64, 56, 56, 64, 3, 1, 1),
    "full_l2": (256, 128, 28, 28, 128, 3, 1, 1),
    "full_l3": (256, 256, 14, 14, 256, 3, 1, 1),
    "full_l4": (256, 512, 7, 7, 512, 3, 1, 1),
    "full_l2_s2": (256, 64, 56, 56, 128, 3, 2, 1),
    "full_l2_ds": (256, 64, 56, 56, 128, 1, 2, 0),
}


def timed(fn, iters=10):
    import torch
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / iters


def err(a, b):
    a, b = a.float(), b.float()
    d = (a - b).abs()
    return {"max_abs": d.max().item(), "rel": (d.max() / (b.abs().max() + 1e-12)).item(),
            "mean_abs": d.mean().item(), "ref_absmax": b.abs().max().item()}


def run_case(name):
    import torch
    import torch.nn.functional as F
    from mpi_pytorch_b200.ops import tc
    torch.manual_seed(0)
    dev = "cuda"
    out = {"case": name}
    if name in GEMM_CASES:
        m, n, k, a_mn, b_mn = GEMM_CASES[name]
        a = torch.randn(m, k, device=dev).bfloat16()
        b = torch.randn(n, k, device=dev).bfloat16()
        ref = a.float() @ b.float().t()
        a_st = a.t().contiguous() if a_mn else a
        b_st = b.t().contiguous() if b_mn else b
        d = tc.gemm(a_st, b_st, m, n, k, bool(a_mn), bool(b_mn), out_dtype=torch.float32)
        torch.cuda.synchronize()
        out["err"] = err(d, ref)
        ms = timed(lambda: tc.gemm(a_st, b_st, m, n, k, bool(a_mn), bool(b_mn)))
        out["ms"] = ms
        out["tflops"] = 2.0 * m * n * k / ms / 1e9
        ms_ref = timed(lambda: a @ b.t())
        out["cublas_ms"] = ms_ref
    else:
        n, c, h, w, k, r, stride, pad = CONV_CASES[name]
        x = torch.randn(n, c, h, w, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
        wt = (torch.randn(k, c, r, r, device=dev) * 0.1).bfloat16().contiguous(memory_format=torch.channels_last)
        xf, wf = x.float().requires_grad_(True), wt.float().requires_grad_(True)
        yref = F.conv2d(xf, wf, None, stride, pad)
        gy = torch.randn_like(yref).bfloat16().contiguous(memory_format=torch.channels_last)
        yref.backward(gy.float())
        st, pd = (stride, stride), (pad, pad)
        y, stats = tc.conv_fprop(x, wt, None, st, pd, True)
        torch.cuda.synchronize()
        out["fprop"] = err(y, yref)
        out["stats_sum"] = err(stats[:k], yref.sum(dim=(0, 2, 3)))
        out["stats_sq"] = err(stats[k:], (yref * yref).sum(dim=(0, 2, 3)))
        flops = 2.0 * y.numel() * c * r * r
        ms = timed(lambda: tc.conv_fprop(x, wt, None, st, pd, False))
        out["fprop_ms"], out["fprop_tflops"] = ms, flops / ms / 1e9
        out["cudnn_fprop_ms"] = timed(lambda: F.conv2d(x, wt, None, stride, pad))
        if tc.dgrad_ok(x, wt, st, pd, (1, 1), 1):  # includes strided
            dx = tc.conv_dgrad(gy, wt, x.shape, st, pd)
            torch.cuda.synchronize()
            out["dgrad"] = err(dx, xf.grad)
            ms = timed(lambda: tc.conv_dgrad(gy, wt, x.shape, st, pd))
            out["dgrad_ms"], out["dgrad_tflops"] = ms, flops / ms / 1e9
        dw = tc.conv_wgrad(gy, x, wt.shape, st, pd)
        torch.cuda.synchronize()
        out["wgrad"] = err(dw, wf.grad)
        ms = timed(lambda: tc.conv_wgrad(gy, x, wt.shape, st, pd))
        out["wgrad_ms"], out["wgrad_tflops"] = ms, flops / ms / 1e9
        out["cudnn_bwd_ms"] = timed(lambda: torch.ops.aten.convolution_backward(
            gy, x, wt, None, [stride, stride], [pad, pad], [1, 1], False, [0, 0], 1, [True, True, False]))
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--case", default=None)
    ap.add_argument("--only", default=None, help="comma list of case-name prefixes")
    ap.add_argument("--timeout", type=int, default=60)
    args = ap.parse_args()
    if args.case:
        run_case(args.case)
        return
    names = list(GEMM_CASES) + list(CONV_CASES)
    if args.only:
        pre = args.only.split(",")
        names = [n for n in names if any(n.startswith(p) for p in pre)]
    os.makedirs("gpurun_out", exist_ok=True)
    with open("gpurun_out/tc_diag.jsonl", "a") as log:
        for n in names:
            try:
                r = subprocess.run([sys.executable, os.path.abspath(__file__), "--case", n], capture_output=True,
                                   text=True, timeout=args.timeout)
                line = r.stdout.strip().splitlines()[-1] if r.stdout.strip() else json.dumps(
                    {"case": n, "rc": r.returncode, "stderr": r.stderr[-600:]})
            except subprocess.TimeoutExpired:
                line = json.dumps({"case": n, "timeout": True})
            print(line, flush=True)
            log.write(line + "\n")


if __name__ == "__main__":
    main()
