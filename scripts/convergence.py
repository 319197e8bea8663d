"""Convergence of the bf16-native B200 path against the UNMODIFIED reference loop on identical data and weights.

The reference's only convergence evidence is a run log: ten epochs over its 800-sample DEBUG set, loss 5.27 -> 1.00,
accuracy on the training sample 0.051 -> 0.986 (/root/reference/training.log:1034-1112; sizes from main.py:75-79).
This script produces the analogue on synthetic data, twice:

  python scripts/convergence.py --arm reference   # CPU fp32: baseline/_ref models.py + mpi_tools.py, main.py:142-185 order
  python scripts/convergence.py --arm ours        # cuda:0, bf16, this framework's kernels and engine (Trainer._step_for)

Both arms see the same 800 images (a fixed low-resolution template per class plus noise, so the labels can be
learnt), the same labels, the same per-epoch permutation and the same initial weights (the reference factory under
``torch.manual_seed(0)``; our model loads that state_dict -- the checkpoint contract makes the names identical).
Each arm writes ``profiles/convergence_<arm>.json`` (per-step loss, per-epoch loss and training-sample accuracy);
``--merge`` joins them into ``profiles/convergence.json``, which tests/test_convergence_gpu.py replays.
"""
import argparse
import json
import os
import sys
import time

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

N_SAMPLES, CLASSES, HW, BATCH, EPOCHS, LR = 800, 100, 128, 128, 10, 4e-4      # main.py:75-79, utils.py:33-42 (DEBUG sizes)
SIGNAL = 0.06          # weight of the class template against the noise: small enough that ten epochs are needed
MEAN, STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)


def make_data(seed: int = 0):
    """[N,3,HW,HW] fp32 (already normalised like main.py:62-65 would) + int64 labels; CPU, deterministic."""
    g = torch.Generator().manual_seed(1000 + seed)
    labels = torch.randint(0, CLASSES, (N_SAMPLES,), generator=g)
    templates = torch.rand(CLASSES, 3, 16, 16, generator=g)
    up = torch.nn.functional.interpolate(templates, size=(HW, HW), mode="nearest")
    img = (SIGNAL * up[labels] + (1.0 - SIGNAL) * torch.rand(N_SAMPLES, 3, HW, HW, generator=g)).clamp_(0, 1)
    mean, std = torch.tensor(MEAN).view(1, 3, 1, 1), torch.tensor(STD).view(1, 3, 1, 1)
    return ((img - mean) / std).contiguous(), labels


def epoch_order(epoch: int) -> torch.Tensor:
    return torch.randperm(N_SAMPLES, generator=torch.Generator().manual_seed(77 + epoch))


def reference_model():
    """The reference factory on CPU under a fixed seed: the common starting point of both arms."""
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    sys.path[:0] = [os.path.join(ROOT, "baseline", "mpi4py_shim"), ref_dir]
    for m in ("utils", "models", "mpi_tools", "helpers"):
        sys.modules.pop(m, None)
    import warnings
    warnings.filterwarnings("ignore")
    import models as rmodels
    import mpi_tools as rtools
    torch.manual_seed(0)
    model, _ = rmodels.initialize_model("resnet18", CLASSES, False, use_pretrained=False)
    for m in ("utils", "models", "mpi_tools", "helpers"):
        sys.modules.pop(m, None)
    del sys.path[:2]
    return model, rtools


def run_reference(max_steps: int):
    torch.set_num_threads(1)                                  # main.py:51
    x, y = make_data()
    model, rtools = reference_model()
    opt = torch.optim.Adam(model.parameters(), lr=LR)         # main.py:125
    rtools.sync_params(model)                                 # main.py:131 (no-op at size 1)
    crit = torch.nn.CrossEntropyLoss()
    steps, epochs = [], []
    t0 = time.time()
    for ep in range(EPOCHS):
        model.train()
        order = epoch_order(ep)
        tr_loss, nb = 0.0, 0
        for b in range(0, N_SAMPLES, BATCH):
            idx = order[b:b + BATCH]
            out = model(x[idx].float())                       # main.py:149
            loss = crit(out, y[idx])
            opt.zero_grad()
            loss.backward()
            rtools.mpi_avg_grads(model)                       # main.py:154
            opt.step()
            tr_loss += loss.detach().item()
            nb += 1
            steps.append(loss.item())
            if max_steps and len(steps) >= max_steps:
                break
        model.eval()
        correct = 0
        with torch.no_grad():                                 # main.py:173-185: accuracy on the training sample
            for b in range(0, N_SAMPLES, BATCH):
                correct += int((model(x[b:b + BATCH]).argmax(1) == y[b:b + BATCH]).sum())
        epochs.append({"epoch": ep, "loss": tr_loss / nb, "acc": correct / N_SAMPLES})
        print("reference epoch", ep, epochs[-1], "t=%.0fs" % (time.time() - t0), flush=True)
        if max_steps and len(steps) >= max_steps:
            break
    return {"arm": "reference (unmodified baseline/_ref, CPU fp32, 1 thread)", "step_loss": steps, "epochs": epochs}


def run_ours(max_steps: int, dtype: str = "bf16"):
    from mpi_pytorch_b200.engine.trainer import Trainer
    from mpi_pytorch_b200.parallel import bootstrap
    from mpi_pytorch_b200.utils.config import Settings
    x, y = make_data()
    ref, _ = reference_model()
    world = bootstrap.init_world("auto")
    s = Settings(MODEL_NAME="resnet18", NUM_CLASSES=CLASSES, BATCH_SIZE=BATCH, HEIGHT=HW, WIDTH=HW, USE_PRETRAINED=False,
                 DTYPE=dtype, LR=LR, SEED=0)
    tr = Trainer(s, world)
    tr.model.load_state_dict(ref.state_dict())                # identical initial weights (fires the shadow refresh)
    tr.engine.refresh_shadow()
    dev = world.device
    xd = x.to(dev, tr.dtype)
    if tr.channels_last:
        xd = xd.contiguous(memory_format=torch.channels_last)
    yd = y.to(dev)
    steps, epochs = [], []
    for ep in range(EPOCHS):
        tr.model.train()
        order = epoch_order(ep).to(dev)
        losses = []
        for b in range(0, N_SAMPLES, BATCH):
            idx = order[b:b + BATCH]
            xb, yb = xd[idx], yd[idx]
            if tr.channels_last:
                xb = xb.contiguous(memory_format=torch.channels_last)
            losses.append(tr._step_for(xb)(xb, yb).float().clone())
            if max_steps and len(steps) + len(losses) >= max_steps:
                break
        ls = [float(v) for v in losses]
        steps += ls
        tr.model.eval()
        correct = 0
        with torch.no_grad():
            for b in range(0, N_SAMPLES, BATCH):
                xb = xd[b:b + BATCH]
                if tr.channels_last:
                    xb = xb.contiguous(memory_format=torch.channels_last)
                correct += int((tr.model(xb).float().argmax(1) == yd[b:b + BATCH]).sum())
        epochs.append({"epoch": ep, "loss": sum(ls) / len(ls), "acc": correct / N_SAMPLES})
        print("ours epoch", ep, epochs[-1], flush=True)
        if max_steps and len(steps) >= max_steps:
            break
    tr.close()
    return {"arm": "ours ({} on {}, native kernels, graph-replayed full batches)".format(dtype, dev),
            "step_loss": steps, "epochs": epochs}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--arm", choices=["reference", "ours"])
    ap.add_argument("--merge", action="store_true")
    ap.add_argument("--max-steps", type=int, default=0)
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles"))
    args = ap.parse_args()
    os.makedirs(args.out, exist_ok=True)
    if args.arm:
        res = run_reference(args.max_steps) if args.arm == "reference" else run_ours(args.max_steps)
        with open(os.path.join(args.out, "convergence_{}.json".format(args.arm)), "w") as fh:
            json.dump(res, fh, indent=1)
    if args.merge:
        a = json.load(open(os.path.join(ROOT, "profiles", "convergence_reference.json")))
        b = json.load(open(os.path.join(args.out, "convergence_ours.json")))
        rows = [{"epoch": r["epoch"], "ref_loss": r["loss"], "ours_loss": o["loss"], "ref_acc": r["acc"], "ours_acc": o["acc"]}
                for r, o in zip(a["epochs"], b["epochs"])]
        n = min(len(a["step_loss"]), len(b["step_loss"]))
        band = max(abs(p - q) / max(p, 0.05) for p, q in zip(a["step_loss"][:n], b["step_loss"][:n]))
        out = {"config": {"model": "resnet18", "samples": N_SAMPLES, "classes": CLASSES, "hw": HW, "batch": BATCH,
                          "epochs": EPOCHS, "lr": LR, "optimizer": "adam"},
               "reference": a, "ours": b, "per_epoch": rows, "max_rel_step_loss_gap": band}
        with open(os.path.join(ROOT, "profiles", "convergence.json"), "w") as fh:
            json.dump(out, fh, indent=1)
        for r in rows:
            print(r)
        print("max relative per-step loss gap:", band)


if __name__ == "__main__":
    main()
