"""Diagnostic: distance between parameter vectors after 3 steps, eager vs eager, graph vs graph, eager vs graph."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mpi_pytorch_b200.engine.trainer import Trainer
from mpi_pytorch_b200.parallel import bootstrap
from mpi_pytorch_b200.utils.config import Settings

def run(graphed, opt, hw=64, batch=16, steps=3):
    w = bootstrap.init_world("cuda")
    s = Settings(MODEL_NAME="resnet18", NUM_CLASSES=16, BATCH_SIZE=batch, HEIGHT=hw, WIDTH=hw, USE_PRETRAINED=False,
                 DTYPE="bf16", LR=1e-2 if opt == "sgd" else 1e-3, SEED=0, OPTIMIZER=opt)
    tr = Trainer(s, w)
    n = tr.engine.layout.total
    p0 = tr.engine.param_arena[:n].clone()
    losses = []
    for x, y in list(tr.make_shard(total=batch * steps, shuffle=False)):
        losses.append(float((tr.train_step_graphed if graphed else tr.train_step)(x, y)))
    torch.cuda.synchronize()
    out = tr.engine.param_arena[:n].clone()
    tr.close()
    return p0, out, losses

for opt in ("sgd", "adam"):
    for hw, batch in ((64, 16), (128, 64)):
        p0, e1, l1 = run(False, opt, hw, batch)
        _, e2, l2 = run(False, opt, hw, batch)
        _, g1, l3 = run(True, opt, hw, batch)
        _, g2, l4 = run(True, opt, hw, batch)
        mv = float((e1 - p0).norm())
        print(opt, hw, batch, "moved %.4f ee %.5f gg %.5f eg %.5f" % (mv, float((e1 - e2).norm()), float((g1 - g2).norm()), float((e1 - g1).norm())),
              "losses", [round(v, 4) for v in l1], [round(v, 4) for v in l3])
