"""CPU tests of host-side helpers behind the CUDA paths: the per-step scratch pool and the stem filter gather maps."""
import importlib.util
import os

import torch
import torch.nn.functional as F

from mpi_pytorch_b200.ops import scratch


def test_scratch_pool_hands_out_disjoint_zeroed_slices_and_falls_back():
    buf = torch.ones(1024)
    scratch.arm(buf)
    try:
        a, pa = scratch.take(100, "cpu")
        b, pb = scratch.take(200, "cpu")
        assert pa and pb and a.numel() == 100 and b.numel() == 200
        assert float(a.abs().sum()) == 0 and float(b.abs().sum()) == 0          # arm() zeroed the pool
        assert a.data_ptr() + 100 * 4 <= b.data_ptr()                           # disjoint, 256-byte aligned
        assert (b.data_ptr() - a.data_ptr()) % 256 == 0
        big, pooled = scratch.take(4096, "cpu")                                 # does not fit: plain zeros
        assert not pooled and big.numel() == 4096 and float(big.abs().sum()) == 0
    finally:
        scratch.disarm()
    c, pc = scratch.take(8, "cpu")
    assert not pc and float(c.abs().sum()) == 0                                 # unarmed: torch.zeros


def _load_tc_index_fn():
    """ops/tc.py binds the CUDA library at import; pull the pure-python map builder out of its source instead."""
    path = os.path.join(os.path.dirname(scratch.__file__), "tc.py")
    src = open(path).read()
    start = src.index("def _s2d_index(")
    end = src.index("def _s2d_geom(")
    ns = {"torch": torch, "_S2D_IDX": {}}
    exec(src[start:end], ns)
    return ns["_s2d_index"]


def test_stem_space_to_depth_filter_maps_reproduce_the_convolution():
    """7x7/2 conv == 4x4/1 conv over the 2x2 space-to-depth input with the gathered filter (all paddings 0..3)."""
    s2d_index = _load_tc_index_fn()
    torch.manual_seed(0)
    c, k = 3, 8
    for pad in range(4):
        x = torch.randn(2, c, 20, 24)
        w = torch.randn(k, c, 7, 7)
        ref = F.conv2d(x, w, None, 2, pad)
        fwd, inv = s2d_index(c, pad, "cpu")
        wk = w.permute(0, 2, 3, 1).reshape(k, 49 * c)
        w2 = torch.cat([wk, wk.new_zeros(k, 1)], 1)[:, fwd].reshape(k, 4, 4, 16)          # [K][a][a'][16]
        assert torch.equal(w2.reshape(k, -1)[:, inv], wk)                                 # inverse map round-trips
        # space-to-depth input with 16-element pixels: [(b*2 + b')*c + ch]
        n, _, h, wd = x.shape
        xs = torch.zeros(n, h // 2, wd // 2, 16)
        for b in range(2):
            for b2 in range(2):
                xs[..., (b * 2 + b2) * c:(b * 2 + b2 + 1) * c] = x[:, :, b::2, b2::2].permute(0, 2, 3, 1)
        lpad = (pad + 1) // 2
        p, q = ref.shape[2], ref.shape[3]
        xs_nchw = F.pad(xs.permute(0, 3, 1, 2), (lpad, 4, lpad, 4))                       # zero halo, generous right/bottom
        out = F.conv2d(xs_nchw, w2.permute(0, 3, 1, 2), None, 1, 0)[:, :, :p, :q]
        assert torch.allclose(out, ref, atol=1e-4, rtol=1e-4)
