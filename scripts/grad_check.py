"""Per-parameter gradient comparison of the native path (our kernels) against the torch/cuDNN path.

Runs one forward+backward (lr = 0, so parameters stay put) of a zoo model on identical synthetic data in both
modes and reports, for every parameter, the relative error of the gradient; mismatching layers show up at once.
  python scripts/grad_check.py --model alexnet --hw 96
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def run(model, hw, batch, dtype, native):
    from mpi_pytorch_b200 import ops
    from mpi_pytorch_b200.engine.trainer import Trainer
    from mpi_pytorch_b200.parallel import bootstrap
    from mpi_pytorch_b200.utils.config import Settings
    ops.set_native(native)
    w = bootstrap.init_world("cuda")
    s = Settings(MODEL_NAME=model, NUM_CLASSES=24, BATCH_SIZE=batch, HEIGHT=hw, WIDTH=hw, USE_PRETRAINED=False,
                 DTYPE=dtype, LR=0.0, SEED=0)
    tr = Trainer(s, w)
    for m in tr.model.modules():                       # make the comparison deterministic
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
    x, y = next(iter(tr.make_shard(total=batch, shuffle=False)))
    loss = tr.forward_loss(x, y)
    tr.optimizer.zero_grad()
    loss.backward()
    torch.cuda.synchronize()
    grads = {n: tr.engine.layout.view(tr.engine.grad_arena, p).detach().float().clone()
             for n, p in tr.model.named_parameters() if p.requires_grad}
    fb = dict(ops.native().FALLBACKS) if native else {}
    tr.optimizer.step()
    tr.close()
    ops.set_native(True)
    return float(loss), grads, fb


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--model", default="resnet18")
    ap.add_argument("--hw", type=int, default=64)
    ap.add_argument("--batch", type=int, default=8)
    ap.add_argument("--dtype", default="bf16")
    args = ap.parse_args()
    l1, g1, fb = run(args.model, args.hw, args.batch, args.dtype, True)
    l0, g0, _ = run(args.model, args.hw, args.batch, args.dtype, False)
    worst = []
    for n in g0:
        a, b = g1[n], g0[n]
        denom = b.abs().max().item() + 1e-12
        worst.append((((a - b).abs().max().item()) / denom, n, a.abs().max().item(), b.abs().max().item()))
    worst.sort(reverse=True)
    print(json.dumps({"model": args.model, "loss_native": l1, "loss_torch": l0, "fallbacks": fb,
                      "worst": [{"rel": round(r, 4), "param": n, "native_absmax": am, "torch_absmax": bm}
                                for r, n, am, bm in worst[:8]],
                      "n_params": len(worst), "n_rel_gt_0.1": sum(1 for r, *_ in worst if r > 0.1)}))


if __name__ == "__main__":
    main()
