"""Data-parallel engine on gloo ranks: DP equivalence, mpi_tools surface, fused optimizer state layout."""
import copy

import pytest
import torch
import torch.nn as nn

from mp_util import run_ranks
from mpi_pytorch_b200.models import initialize_model
from mpi_pytorch_b200.parallel import bootstrap
from mpi_pytorch_b200.parallel.arena import FlatLayout
from mpi_pytorch_b200.parallel.collectives import Hyper
from mpi_pytorch_b200.parallel.ddp import FusedDataParallel


def _tiny():
    torch.manual_seed(0)
    return nn.Sequential(nn.Conv2d(3, 8, 3, padding=1, bias=False), nn.BatchNorm2d(8), nn.ReLU(), nn.AdaptiveAvgPool2d(1),
                         nn.Flatten(), nn.Linear(8, 5))


def test_layout_buckets_reverse_order_and_alignment():
    m = _tiny()
    lay = FlatLayout(list(m.parameters()), bucket_bytes=1024, world=4)
    assert lay.total % (64 * 4) == 0
    first = lay.buckets[0].params[0]
    assert first is list(m.parameters())[-1]                 # head first == autograd order
    for b in lay.buckets:
        assert b.start % 64 == 0 and b.numel % (64 * 4) == 0
    offs = sorted((s.offset, s.numel) for s in lay.slots.values())
    for (o1, n1), (o2, _) in zip(offs, offs[1:]):
        assert o1 + n1 <= o2


def test_single_process_engine_matches_torch_adam():
    torch.manual_seed(1)
    ref = _tiny()
    mine = copy.deepcopy(ref)
    opt = torch.optim.Adam(ref.parameters(), lr=1e-2)
    w = bootstrap.init_world("cpu")
    eng = FusedDataParallel(mine, w, Hyper("adam", lr=1e-2), bucket_mb=0.001)
    for _ in range(3):
        x, y = torch.randn(4, 3, 8, 8), torch.randint(0, 5, (4,))
        l1 = nn.functional.cross_entropy(ref(x), y)
        opt.zero_grad(); l1.backward(); opt.step()
        l2 = nn.functional.cross_entropy(mine(x), y)
        eng.optimizer.zero_grad(); l2.backward(); eng.optimizer.step()
        assert torch.allclose(l1, l2, atol=1e-6)
    for a, b in zip(ref.parameters(), mine.parameters()):
        assert torch.allclose(a, b, atol=1e-6)
    sd_ref, sd = opt.state_dict(), eng.optimizer.state_dict()
    assert sd["param_groups"][0]["params"] == sd_ref["param_groups"][0]["params"]
    for k in sd_ref["state"]:
        assert torch.allclose(sd["state"][k]["exp_avg"], sd_ref["state"][k]["exp_avg"], atol=1e-7)
        assert float(sd["state"][k]["step"]) == float(sd_ref["state"][k]["step"])
    # a stock Adam accepts our state_dict (the reference's load_checkpoint path, helpers.py:14)
    stock = torch.optim.Adam(copy.deepcopy(ref).parameters(), lr=1e-2)
    stock.load_state_dict(sd)
    # and ours accepts a stock one
    eng.optimizer.load_state_dict(sd_ref)
    assert eng.step_count == 3


def _tiny_nobn():
    torch.manual_seed(0)
    return nn.Sequential(nn.Conv2d(3, 8, 3, padding=1), nn.ReLU(), nn.AdaptiveAvgPool2d(1), nn.Flatten(), nn.Linear(8, 5))


def _dp_worker(w, steps, bn):
    torch.manual_seed(0)
    model = _tiny() if bn else _tiny_nobn()
    eng = FusedDataParallel(model, w, Hyper("adam", lr=1e-2), bucket_mb=0.001)
    if w.rank != 0:                                   # de-synchronise, then sync_params must repair it
        with torch.no_grad():
            for p in model.parameters():
                p.add_(1.0)
            if bn:
                model[1].running_mean.add_(3.0)
    eng.sync_params(0)
    synced_rm = model[1].running_mean.clone() if bn else torch.zeros(1)
    g = torch.Generator().manual_seed(123)
    X, Y = torch.randn(steps, 8, 3, 8, 8, generator=g), torch.randint(0, 5, (steps, 8), generator=g)
    per = 8 // w.size
    for s in range(steps):
        x, y = X[s, w.rank * per:(w.rank + 1) * per], Y[s, w.rank * per:(w.rank + 1) * per]
        loss = nn.functional.cross_entropy(model(x), y)
        eng.optimizer.zero_grad(); loss.backward(); eng.optimizer.step()
    return [p.detach().clone() for p in model.parameters()], synced_rm


def test_dp_equivalence_2_ranks_vs_1():
    """N ranks x batch B == 1 rank x batch N*B (mpi_tools.py:30-37 semantics: SUM then / size).  BatchNorm uses
    per-rank statistics, so the equality with the single-process run is checked on a BN-free model; the BN model
    checks replica agreement and that sync_params repaired the de-synchronised buffers."""
    two = run_ranks(_dp_worker, 2, (3, False))
    one = run_ranks(_dp_worker, 1, (3, False))
    for a, b in zip(two[0][0], two[1][0]):
        assert torch.equal(a, b)                       # replicas stay bit-identical
    for a, ref in zip(two[0][0], one[0][0]):
        assert torch.allclose(a, ref, atol=2e-6, rtol=1e-5), float((a - ref).abs().max())
    res = run_ranks(_dp_worker, 2, (3, True))
    for a, b in zip(res[0][0], res[1][0]):
        assert torch.equal(a, b)
    assert torch.equal(res[0][1], res[1][1])           # rank 1's +3.0 running_mean was overwritten by rank 0's
    assert float(res[1][1].abs().max()) < 1.0


class _OddBuffers(nn.Module):
    """Parameter + buffer element counts that are NOT multiples of 4 (DenseNet-121: 83,769 buffer elements)."""

    def __init__(self):
        super().__init__()
        self.w = nn.Parameter(torch.arange(7, dtype=torch.float32))
        self.frozen = nn.Parameter(torch.arange(5, dtype=torch.float32), requires_grad=False)
        self.register_buffer("stat", torch.arange(9, dtype=torch.float32))
        self.register_buffer("count", torch.tensor(3, dtype=torch.long))


def _odd_sync_worker(w):
    from mpi_pytorch_b200.parallel import mpi_tools
    out = []
    for front in ("engine", "mpi_tools"):
        m = _OddBuffers()
        if w.rank != 0:
            with torch.no_grad():
                m.w.add_(10.0); m.frozen.add_(10.0); m.stat.add_(10.0); m.count.add_(10)
        if front == "engine":
            FusedDataParallel(m, w, Hyper("adam")).sync_params(0)
        else:
            mpi_tools.sync_params(m)
        out.append((m.w.detach().clone(), m.frozen.detach().clone(), m.stat.clone(), m.count.clone()))
    return out


def test_sync_params_odd_sizes_2_ranks():
    res = run_ranks(_odd_sync_worker, 2)
    for front in range(2):
        for a, b in zip(res[0][front], res[1][front]):
            assert torch.equal(a, b)
        assert torch.equal(res[1][front][0], torch.arange(7, dtype=torch.float32))
        assert int(res[1][front][3]) == 3


def _avg_worker(w):
    from mpi_pytorch_b200.parallel import mpi_tools
    torch.manual_seed(0)
    model = _tiny()
    for p in model[0].parameters():
        p.requires_grad_(False)                        # frozen params must not crash
    with torch.no_grad():
        if w.rank == 1:
            for p in model.parameters():
                p.mul_(2.0)
    mpi_tools.sync_params(model)
    x = torch.full((2, 3, 8, 8), float(w.rank + 1))
    model(x).sum().backward()
    local = [p.grad.clone() for p in model.parameters() if p.requires_grad]
    mpi_tools.mpi_avg_grads(model)
    avg = [p.grad.clone() for p in model.parameters() if p.requires_grad]
    s = mpi_tools.mpi_sum(float(w.rank + 1), mpi_tools.SUM)
    return local, avg, [p.detach().clone() for p in model.parameters()], s, mpi_tools.num_processes()


def test_mpi_tools_surface_2_ranks():
    res = run_ranks(_avg_worker, 2)
    (l0, a0, p0, s0, n0), (l1, a1, p1, s1, n1) = res
    assert n0 == n1 == 2 and s0 == s1 == 3.0
    for x, y in zip(p0, p1):
        assert torch.equal(x, y)
    for g0, g1, m0, m1 in zip(l0, l1, a0, a1):
        assert torch.allclose(m0, (g0 + g1) / 2, atol=1e-6)
        assert torch.equal(m0, m1)


def _avg_bf16_worker(w):
    from mpi_pytorch_b200.parallel import mpi_tools
    torch.manual_seed(0)
    model = _tiny().to(torch.bfloat16)
    mpi_tools.sync_params(model)
    x = torch.full((2, 3, 8, 8), float(w.rank + 1), dtype=torch.bfloat16)
    model(x).sum().backward()
    local = [p.grad.clone() for p in model.parameters()]
    mpi_tools.mpi_avg_grads(model)
    return local, [p.grad.clone() for p in model.parameters()], mpi_tools._AVERAGERS[model].dtype


def test_mpi_avg_grads_bf16_wire_2_ranks():
    """An all-bf16 model exchanges its gradients in bf16 (2 bytes per element on the wire)."""
    (l0, a0, d0), (l1, a1, d1) = run_ranks(_avg_bf16_worker, 2)
    assert d0 == d1 == torch.bfloat16
    for g0, g1, m0, m1 in zip(l0, l1, a0, a1):
        assert m0.dtype == torch.bfloat16 and torch.equal(m0, m1)
        assert torch.allclose(m0.float(), (g0.float() + g1.float()) / 2, rtol=2e-2, atol=1e-3)


def test_mpi_tools_single_process_noop():
    from mpi_pytorch_b200.parallel import mpi_tools
    m = _tiny()
    assert mpi_tools.num_processes() == 1
    assert mpi_tools.mpi_avg_grads(m) is None and mpi_tools.sync_params(m) is None
    assert mpi_tools.mpi_sum(2.5, mpi_tools.SUM) == 2.5


def test_bucket_schedule_ends_small():
    """The last buckets to fire in backward cannot overlap with compute, so they are the small ones."""
    m, _ = initialize_model("resnet34", 1000, False, use_pretrained=False)
    lay = FlatLayout(list(m.parameters()), bucket_bytes=32 << 20, world=8)
    mb = [b.numel * 4 / 2 ** 20 for b in lay.buckets]
    assert mb[-1] <= 2.0 + 1e-6 and mb[-2] <= 8.0 + 1e-6 and max(mb) <= 32.0 + 1e-6
    assert sum(len(b.params) for b in lay.buckets) == len(lay.params)
    offs = sorted((s.offset, s.numel) for s in lay.slots.values())
    for (o1, n1), (o2, _) in zip(offs, offs[1:]):
        assert o1 + n1 <= o2
    flat = FlatLayout(list(m.parameters()), bucket_bytes=32 << 20, world=8, tail_bytes=32 << 20)
    assert len(flat.buckets) < len(lay.buckets)


def _wire_worker(w, wire):
    torch.manual_seed(0)
    model = _tiny_nobn()
    eng = FusedDataParallel(model, w, Hyper("adam", lr=1e-2), bucket_mb=0.001,
                            grad_dtype=torch.bfloat16 if wire else torch.float32)
    eng.sync_params(0)
    g = torch.Generator().manual_seed(5)
    X, Y = torch.randn(2, 8, 3, 8, 8, generator=g), torch.randint(0, 5, (2, 8), generator=g)
    per = 8 // w.size
    for s in range(2):
        x, y = X[s, w.rank * per:(w.rank + 1) * per], Y[s, w.rank * per:(w.rank + 1) * per]
        loss = nn.functional.cross_entropy(model(x), y)
        eng.optimizer.zero_grad(); loss.backward(); eng.optimizer.step()
    return [p.detach().clone() for p in model.parameters()], eng.wire_arena is not None


def test_bf16_wire_format_in_the_engine_2_ranks():
    """GRAD_DTYPE=bf16: gradients cross the wire as bf16 (one rounding per rank), sums and Adam stay fp32."""
    a = run_ranks(_wire_worker, 2, (True,))
    b = run_ranks(_wire_worker, 2, (False,))
    assert a[0][1] and not b[0][1]
    for x, y in zip(a[0][0], a[1][0]):
        assert torch.equal(x, y)
    for x, y in zip(a[0][0], b[0][0]):
        assert torch.allclose(x, y, atol=2.5e-2)          # Adam's normalised step: bounded by ~2 * lr per step
        assert not torch.equal(x, y) or x.numel() < 4


def test_engine_suspended_keeps_gradients_local():
    w = bootstrap.init_world("cpu")
    torch.manual_seed(0)
    model = _tiny_nobn()
    eng = FusedDataParallel(model, w, Hyper("adam", lr=1e-2))
    before = [p.detach().clone() for p in model.parameters()]
    with eng.suspended():
        nn.functional.cross_entropy(model(torch.randn(4, 3, 8, 8)), torch.randint(0, 5, (4,))).backward()
    assert all(torch.equal(a, b) for a, b in zip(before, model.parameters()))
    assert float(eng.grad_arena.abs().sum()) > 0 and eng.step_count == 0
