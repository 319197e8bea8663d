"""Eval pipeline: summed accuracy over N staged ranks == single-process accuracy (SURVEY §4 plan)."""
import torch
import torch.nn as nn

from mp_util import run_ranks
from mpi_pytorch_b200.parallel import bootstrap
from mpi_pytorch_b200.parallel.pipeline import EvalPipeline, assign_roles

N, RAW, OUT, NCLS = 37, (20, 14), (8, 8), 4


def _source(i0, i1):
    g = torch.Generator().manual_seed(1000 + i0)
    return torch.randint(0, 256, (i1 - i0, RAW[0], RAW[1], 3), generator=g, dtype=torch.uint8)


def _model_fn():
    torch.manual_seed(5)
    return nn.Sequential(nn.Conv2d(3, 4, 3, padding=1), nn.ReLU(), nn.AdaptiveAvgPool2d(1), nn.Flatten(),
                         nn.Linear(4, NCLS))


def _labels():
    m = _model_fn().eval()
    from mpi_pytorch_b200 import ops
    outs = []
    with torch.no_grad():
        for i0 in range(0, N, 5):
            x = ops.normalize_u8(ops.resize_u8(_source_fixed(i0, min(i0 + 5, N)), OUT[0], OUT[1]))
            outs.append(m(x).argmax(1))
    pred = torch.cat(outs)
    lab = pred.clone()
    lab[::3] = (lab[::3] + 1) % NCLS                       # make ~1/3 wrong so accuracy is non-trivial
    return lab


def _source_fixed(i0, i1):
    # per-sample deterministic frames, independent of micro-batch boundaries
    return torch.stack([_source(i, i + 1)[0] for i in range(i0, i1)])


def _worker(w, labels):
    from mpi_pytorch_b200.parallel.collectives import make_collectives
    coll = make_collectives(w) if w.size > 2 else None      # also cover the collectives-object count reduction (K13)
    pipe = EvalPipeline(w, N, RAW, OUT, _source_fixed, labels, _model_fn, micro_batch=5, seed=3, collectives=coll)
    return pipe.run(), int(pipe.seen)


def test_roles():
    r = assign_roles(8)
    assert (r.reader, r.resizer, r.normalizer, r.predictors) == (0, 1, 2, [3, 4, 5, 6, 7])
    assert assign_roles(1).predictors == [0] and assign_roles(2).predictors == [1]


def test_pipeline_4_ranks_equals_single():
    labels = _labels()
    expected = float((torch.arange(N) % 3 != 0).sum()) / N
    w = bootstrap.init_world("cpu")
    acc1, seen1 = _worker(w, labels)
    assert seen1 == N and abs(acc1 - expected) < 1e-9
    for n in (2, 4, 5):
        res = run_ranks(_worker, n, (labels,))
        assert abs(res[0][0] - expected) < 1e-9
        assert sum(s for _, s in res) == N              # every image predicted exactly once


def _source_ragged(i0, i1):
    # frames of two different true sizes inside one canvas-sized message
    out = []
    for i in range(i0, i1):
        f = _source(i, i + 1)[0]
        out.append(f if i % 2 == 0 else f[:RAW[0] - 4, :RAW[1] - 3].contiguous())
    return out


def _ragged_worker(w, labels, hw):
    pipe = EvalPipeline(w, N, RAW, OUT, _source_ragged, labels, _model_fn, micro_batch=5, seed=3, frame_hw=hw)
    return pipe.run(), int(pipe.seen), pipe.stats


def test_pipeline_resizes_each_frame_over_its_own_extent():
    """A small image inside a larger canvas must be resized over ITS size (evaluation_pipeline.py:89 resizes each PIL
    image individually), not squeezed into the canvas corner: predictions equal per-image preprocessing."""
    from mpi_pytorch_b200 import ops
    m = _model_fn().eval()
    hw = torch.tensor([[RAW[0], RAW[1]] if i % 2 == 0 else [RAW[0] - 4, RAW[1] - 3] for i in range(N)])
    with torch.no_grad():
        labels = torch.cat([m(ops.normalize_u8(ops.resize_u8(f[None], OUT[0], OUT[1]))).argmax(1)
                            for f in _source_ragged(0, N)])
    labels[::4] = (labels[::4] + 1) % NCLS
    expected = float((torch.arange(N) % 4 != 0).sum()) / N
    res = run_ranks(_ragged_worker, 4, (labels, hw))
    assert abs(res[0][0] - expected) < 1e-9 and sum(s for _, s, _ in res) == N
    assert res[0][2]["images_per_s"] > 0
