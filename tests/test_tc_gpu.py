"""tcgen05 GEMM / implicit-GEMM convolution vs an fp32 PyTorch oracle on the same bf16-rounded inputs."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def _tc():
    from mpi_pytorch_b200.ops import tc
    return tc


def _rel(a, b):
    return ((a.float() - b.float()).abs().max() / (b.float().abs().max() + 1e-12)).item()


@pytest.mark.parametrize("m,n,k,a_mn,b_mn", [(128, 64, 64, 0, 0), (256, 1000, 512, 0, 0), (300, 136, 200, 0, 0),
                                              (256, 512, 1000, 0, 1), (1000, 512, 256, 1, 1), (512, 256, 768, 1, 0)])
def test_gemm_all_majorness(m, n, k, a_mn, b_mn):
    torch.manual_seed(0)
    a = torch.randn(m, k, device="cuda").bfloat16()
    b = torch.randn(n, k, device="cuda").bfloat16()
    ref = a.float() @ b.float().t()
    a_st = a.t().contiguous() if a_mn else a
    b_st = b.t().contiguous() if b_mn else b
    d32 = _tc().gemm(a_st, b_st, m, n, k, bool(a_mn), bool(b_mn), out_dtype=torch.float32)
    assert _rel(d32, ref) < 1e-5
    bias = torch.randn(n, device="cuda")
    d16 = _tc().gemm(a_st, b_st, m, n, k, bool(a_mn), bool(b_mn), bias=bias)
    assert _rel(d16, ref + bias) < 1e-2


def test_gemm_splitk_atomic():
    a = torch.randn(128, 4096, device="cuda").bfloat16()
    b = torch.randn(64, 4096, device="cuda").bfloat16()
    d = _tc().gemm(a, b, 128, 64, 4096, False, False, out_dtype=torch.float32, splitk=8)
    assert _rel(d, a.float() @ b.float().t()) < 1e-5


CONVS = [(2, 64, 8, 8, 64, 1, 1, 0), (2, 64, 8, 8, 64, 3, 1, 1), (8, 64, 56, 56, 64, 3, 1, 1),
         (8, 64, 56, 56, 128, 3, 2, 1), (8, 64, 56, 56, 128, 1, 2, 0), (8, 256, 14, 14, 256, 3, 1, 1),
         (16, 512, 7, 7, 512, 3, 1, 1), (4, 64, 17, 17, 128, 3, 1, 0), (3, 128, 9, 11, 192, 5, 1, 2),
         # channel counts that are not multiples of 64 (DenseNet / Inception / SqueezeNet): TMA zero-fills the tails
         (4, 96, 12, 12, 48, 3, 1, 1), (4, 16, 13, 13, 64, 1, 1, 0), (2, 160, 9, 9, 192, 3, 1, 1),
         (2, 288, 17, 17, 384, 3, 2, 0), (2, 80, 10, 10, 192, 3, 1, 0),
         # more strided shapes: odd sizes, 5x5 stride 2, stride 3, 1x1 stride 2 on odd maps
         (2, 64, 15, 15, 64, 5, 2, 2), (2, 64, 20, 20, 64, 3, 3, 1), (2, 64, 9, 9, 128, 1, 2, 0), (2, 64, 14, 14, 64, 7, 2, 3),
         # halo-tile wgrad: row mode with one / several channel and k blocks, wide and short maps, many images per CTA
         (2, 128, 10, 10, 128, 3, 1, 1), (3, 128, 6, 40, 64, 3, 1, 1), (40, 256, 7, 7, 128, 3, 1, 1),
         (37, 64, 28, 28, 64, 3, 1, 1)]


@pytest.mark.parametrize("n,hw,k,pad", [(4, 64, 64, 3), (3, 38, 128, 2), (2, 30, 64, 0), (5, 32, 64, 1)])
def test_stem_space_to_depth_conv(n, hw, k, pad):
    """7x7/2 stems over 3 channels as a 4x1 convolution over overlapping space-to-depth super-pixels."""
    tc = _tc()
    torch.manual_seed(3)
    x = torch.randn(n, 3, hw, hw + 2, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    wt = (torch.randn(k, 3, 7, 7, device="cuda") * 0.1).bfloat16().contiguous(memory_format=torch.channels_last)
    assert tc.stem_s2d_ok(x, wt, (2, 2), (pad, pad), (1, 1), 1)
    xf, wf = x.float(), wt.float().requires_grad_(True)
    yref = F.conv2d(xf, wf, None, 2, pad)
    gy = torch.randn_like(yref).bfloat16().contiguous(memory_format=torch.channels_last)
    yref.backward(gy.float())
    y, stats, xs = tc.conv_fprop_stem(x, wt, None, (pad, pad), True)
    assert y.shape == yref.shape and _rel(y, yref) < 8e-3
    assert _rel(stats[:k], yref.sum(dim=(0, 2, 3))) < 1e-4
    assert _rel(stats[k:], (yref * yref).sum(dim=(0, 2, 3))) < 1e-4
    dw = tc.conv_wgrad_stem(gy, xs, x.shape, wt.shape, (pad, pad))
    assert dw.shape == wf.grad.shape and _rel(dw, wf.grad) < 2e-4


def test_wgrad_halo_kernel_covers_the_resnet_layers():
    tc = _tc()
    for (c, k, hw) in ((64, 64, 56), (128, 128, 28), (256, 256, 14)):
        assert tc.wgrad_halo_ok(c, k, 3, 3, hw, hw, (1, 1))
    assert not tc.wgrad_halo_ok(512, 512, 3, 3, 7, 7, (1, 1))       # 48 column groups: split-K form is faster
    assert not tc.wgrad_halo_ok(64, 128, 3, 3, 28, 28, (2, 2))      # strided: im2col-mode kernel
    assert not tc.wgrad_halo_ok(64, 64, 1, 1, 56, 56, (1, 1))       # 1x1: plain GEMM
    assert not tc.wgrad_halo_ok(96, 48, 3, 3, 12, 12, (1, 1))       # ragged channel counts: im2col-mode kernel


def test_conv_rect_kernels_inception():
    """1x7 / 7x1 convolutions with asymmetric padding (Inception-C)."""
    tc = _tc()
    for (kh, kw, ph, pw) in ((1, 7, 0, 3), (7, 1, 3, 0)):
        x = torch.randn(2, 128, 17, 17, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
        wt = (torch.randn(160, 128, kh, kw, device="cuda") * 0.1).bfloat16().contiguous(memory_format=torch.channels_last)
        xf, wf = x.float().requires_grad_(True), wt.float().requires_grad_(True)
        yref = F.conv2d(xf, wf, None, 1, (ph, pw))
        gy = torch.randn_like(yref).bfloat16().contiguous(memory_format=torch.channels_last)
        yref.backward(gy.float())
        y, _ = tc.conv_fprop(x, wt, None, (1, 1), (ph, pw), False)
        assert _rel(y, yref) < 8e-3
        assert _rel(tc.conv_dgrad(gy, wt, x.shape, (1, 1), (ph, pw)), xf.grad) < 8e-3
        assert _rel(tc.conv_wgrad(gy, x, wt.shape, (1, 1), (ph, pw)), wf.grad) < 2e-4


@pytest.mark.parametrize("n,c,h,w,k,r,stride,pad", CONVS)
def test_conv_fprop_dgrad_wgrad(n, c, h, w, k, r, stride, pad):
    tc = _tc()
    torch.manual_seed(1)
    x = torch.randn(n, c, h, w, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    wt = (torch.randn(k, c, r, r, device="cuda") * 0.1).bfloat16().contiguous(memory_format=torch.channels_last)
    xf, wf = x.float().requires_grad_(True), wt.float().requires_grad_(True)
    yref = F.conv2d(xf, wf, None, stride, pad)
    gy = torch.randn_like(yref).bfloat16().contiguous(memory_format=torch.channels_last)
    yref.backward(gy.float())
    st, pd = (stride, stride), (pad, pad)
    assert tc.fprop_ok(x, wt, st, pd, (1, 1), 1)
    y, stats = tc.conv_fprop(x, wt, None, st, pd, True)
    assert _rel(y, yref) < 8e-3                                 # bf16 output rounding
    assert _rel(stats[:k], yref.sum(dim=(0, 2, 3))) < 1e-4      # stats come from the fp32 accumulator
    assert _rel(stats[k:], (yref * yref).sum(dim=(0, 2, 3))) < 1e-4
    assert tc.dgrad_ok(x, wt, st, pd, (1, 1), 1)                # stride 1 and strided (parity-decomposed) dgrad
    assert _rel(tc.conv_dgrad(gy, wt, x.shape, st, pd), xf.grad) < 8e-3
    dw = tc.conv_wgrad(gy, x, wt.shape, st, pd)
    assert dw.shape == wf.grad.shape and _rel(dw, wf.grad) < 2e-4
    acc = torch.ones(k, r, r, c, device="cuda")                 # accumulate-into-arena form
    tc.conv_wgrad(gy, x, wt.shape, st, pd, out=acc)
    assert _rel(acc.permute(0, 3, 1, 2), wf.grad + 1.0) < 2e-4


def test_resnet34_bf16_step_uses_tensor_core_kernels():
    """bf16 ResNet-18 steps: native path (tcgen05 convs) tracks the torch/cuDNN autocast path."""
    from mpi_pytorch_b200 import ops
    from mpi_pytorch_b200.engine.trainer import Trainer
    from mpi_pytorch_b200.parallel import bootstrap
    from mpi_pytorch_b200.utils.config import Settings
    w = bootstrap.init_world("cuda")
    s = Settings(MODEL_NAME="resnet18", NUM_CLASSES=32, BATCH_SIZE=16, HEIGHT=64, WIDTH=64, USE_PRETRAINED=False,
                 DTYPE="bf16", LR=1e-3)
    losses = {}
    before = _tc().LAUNCHES["n"]
    for native in (True, False):
        ops.set_native(native)
        try:
            tr = Trainer(s, w)
            shard = tr.make_shard(total=64, shuffle=False)
            losses[native] = [float(tr.train_step(x, y)) for x, y in shard]
            tr.close()
        finally:
            ops.set_native(True)
    assert _tc().LAUNCHES["n"] - before > 100                   # the convs really ran on the tcgen05 kernel
    assert abs(losses[True][0] - losses[False][0]) < 0.05
    assert losses[True][-1] < losses[True][0] + 0.5 and all(l == l for l in losses[True])


@pytest.mark.parametrize("name,hw", [("alexnet", 96), ("vgg", 64), ("squeezenet", 96), ("densenet", 64), ("inception", 128),
                                     ("resnet34", 64)])
def test_every_zoo_member_trains_on_the_native_path(name, hw):
    """One bf16 training step of each architecture through the native ops (finite loss, params move)."""
    from mpi_pytorch_b200.engine.trainer import Trainer
    from mpi_pytorch_b200.parallel import bootstrap
    from mpi_pytorch_b200.utils.config import Settings
    w = bootstrap.init_world("cuda")
    s = Settings(MODEL_NAME=name, NUM_CLASSES=24, BATCH_SIZE=8, HEIGHT=hw, WIDTH=hw, USE_PRETRAINED=False, DTYPE="bf16",
                 LR=1e-3)
    tr = Trainer(s, w)
    before = {n: p.detach().clone() for n, p in tr.model.named_parameters()}
    shard = tr.make_shard(total=16, shuffle=False)
    losses = [float(tr.train_step(x, y)) for x, y in shard]
    assert all(l == l and l < 100 for l in losses), losses
    # every parameter that received a gradient must have been updated by the bucket it lives in (a bucket that
    # fires before its last gradient landed would leave late parameters untouched)
    stale = [n for n, p in tr.model.named_parameters()
             if torch.equal(before[n], p.detach()) and float(tr.engine.layout.view(tr.engine.grad_arena, p).abs().max()) > 0]
    assert not stale, stale
    tr.close()


@pytest.mark.parametrize("n,c,h,w,k,r,stride,pad", [(4, 3, 64, 64, 64, 7, 2, 3), (2, 3, 35, 35, 64, 11, 4, 2),
                                                    (2, 3, 33, 33, 32, 3, 2, 0)])
def test_stem_conv_via_im2col_gemm(n, c, h, w, k, r, stride, pad):
    """3-channel stems: explicit im2col + tensor-core GEMM for fprop and wgrad."""
    tc = _tc()
    torch.manual_seed(2)
    x = torch.randn(n, c, h, w, device="cuda").bfloat16().contiguous(memory_format=torch.channels_last)
    wt = (torch.randn(k, c, r, r, device="cuda") * 0.1).bfloat16().contiguous(memory_format=torch.channels_last)
    xf, wf = x.float(), wt.float().requires_grad_(True)
    yref = F.conv2d(xf, wf, None, stride, pad)
    gy = torch.randn_like(yref).bfloat16().contiguous(memory_format=torch.channels_last)
    yref.backward(gy.float())
    st, pd = (stride, stride), (pad, pad)
    assert tc.im2col_ok(x, wt, st, pd, (1, 1), 1) and not tc.fprop_ok(x, wt, st, pd, (1, 1), 1)
    y, stats, xcol = tc.conv_fprop_im2col(x, wt, None, st, pd, True)
    assert y.shape == yref.shape and _rel(y, yref) < 8e-3
    assert _rel(stats[:k], yref.sum(dim=(0, 2, 3))) < 1e-4
    dw = tc.conv_wgrad_im2col(gy, xcol, wt.shape)
    assert dw.shape == wf.grad.shape and _rel(dw, wf.grad) < 2e-4


def test_linear_head_with_64500_classes_stays_on_the_tensor_core_gemm():
    """The reference's default head (utils.py:39: 64,500 classes, not a multiple of 8) must not fall back to cuBLAS."""
    from mpi_pytorch_b200 import ops
    nat = ops.native()
    torch.manual_seed(0)
    x = (torch.randn(64, 512, device="cuda") * 0.5).to(torch.bfloat16).requires_grad_(True)
    w = (torch.randn(64500, 512, device="cuda") * 0.05).requires_grad_(True)
    b = torch.randn(64500, device="cuda").requires_grad_(True)
    before = dict(nat.FALLBACKS)
    y = nat.linear(x, w, b)
    gy = torch.randn(64, 64500, device="cuda").to(torch.bfloat16)
    y.backward(gy)
    assert dict(nat.FALLBACKS) == before
    xr, wr = x.detach().float().requires_grad_(True), w.detach().to(torch.bfloat16).float().requires_grad_(True)
    yr = torch.nn.functional.linear(xr, wr, b.detach())
    yr.backward(gy.float())
    assert y.shape == (64, 64500) and torch.allclose(y.float(), yr, atol=0.06, rtol=0.03)
    assert torch.allclose(x.grad.float(), xr.grad, atol=0.5, rtol=0.03)
    assert torch.allclose(w.grad, wr.grad, atol=0.15, rtol=0.03)
    assert torch.allclose(b.grad, gy.float().sum(0), atol=0.1, rtol=0.02)
