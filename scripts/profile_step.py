"""Kernel-time table of eager training steps (torch.profiler / CUPTI): where a ResNet-34 step spends its GPU time."""
import collections
import json
import os
import re
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_pytorch_b200.engine.trainer import Trainer  # noqa: E402
from mpi_pytorch_b200.parallel import bootstrap  # noqa: E402
from mpi_pytorch_b200.utils.config import Settings  # noqa: E402


def main():
    model = sys.argv[1] if len(sys.argv) > 1 else "resnet34"
    steps = 3
    w = bootstrap.init_world("cuda")
    bs, hw = int(os.environ.get("MPB_PROF_BATCH", "256")), int(os.environ.get("MPB_PROF_IMAGE", "224"))
    s = Settings(MODEL_NAME=model, NUM_CLASSES=int(os.environ.get("MPB_PROF_CLASSES", "1000")), BATCH_SIZE=bs, HEIGHT=hw, WIDTH=hw, USE_PRETRAINED=False,
                 DTYPE="bf16", LR=4e-4)
    tr = Trainer(s, w)
    batches = list(iter(tr.make_shard(total=2 * bs, shuffle=False)))
    for i in range(3):
        tr.train_step(*batches[i % 2])
    torch.cuda.synchronize()
    from torch.profiler import ProfilerActivity, profile
    want_aten = os.environ.get("MPB_PROF_ATEN", "0") == "1"
    acts = [ProfilerActivity.CUDA] + ([ProfilerActivity.CPU] if want_aten else [])
    with profile(activities=acts, record_shapes=want_aten) as prof:
        for i in range(steps):
            tr.train_step(*batches[i % 2])
        torch.cuda.synchronize()
    if want_aten:      # which framework ops still launch library elementwise kernels (targets for fusion)
        ops = collections.Counter()
        for ev in prof.events():
            if ev.device_type == torch.autograd.DeviceType.CPU and ev.name.startswith("aten::"):
                ops[(ev.name, str(ev.input_shapes)[:80])] += 1
        for (name, shp), n in ops.most_common(40):
            print("aten {:6.1f}x  {:28s} {}".format(n / steps, name, shp))
    agg = collections.defaultdict(lambda: [0, 0.0])
    for ev in prof.events():
        if ev.device_type == torch.autograd.DeviceType.CUDA:
            name = ev.name.replace("(anonymous namespace)::", "").replace("void ", "")
            name = re.sub(r"[<(].*", "", name)[:64]
            agg[name][0] += 1
            agg[name][1] += ev.device_time if hasattr(ev, "device_time") else ev.cuda_time
    total = sum(v[1] for v in agg.values())
    rows = sorted(agg.items(), key=lambda kv: -kv[1][1])
    out = {"model": model, "steps": steps, "gpu_ms_per_step": total / steps / 1e3,
           "kernels": [{"name": k, "launches_per_step": n / steps, "ms_per_step": t / steps / 1e3,
                        "share": t / total} for k, (n, t) in rows[:24]]}
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump(out, open("gpurun_out/profile_step_{}.json".format(model), "w"), indent=1)
    print("gpu ms/step (sum of kernels): {:.2f}".format(out["gpu_ms_per_step"]))
    for r in out["kernels"]:
        print("{:8.3f} ms {:6.1f}x {:5.1f}%  {}".format(r["ms_per_step"], r["launches_per_step"], 100 * r["share"], r["name"]))


if __name__ == "__main__":
    main()
