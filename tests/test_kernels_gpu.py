"""Numerics of every non-GEMM sm_100a kernel against a plain PyTorch fp32 oracle (SURVEY §4 plan)."""
import copy

import pytest
import torch
import torch.nn as nn
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def nat():
    from mpi_pytorch_b200 import ops
    return ops.native()


def cl(x):
    return x.contiguous(memory_format=torch.channels_last)


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 2e-5), (torch.bfloat16, 2e-2)])
@pytest.mark.parametrize("C", [1000, 64500, 37])
def test_softmax_ce(dtype, tol, C):
    torch.manual_seed(0)
    B = 32
    logits = (torch.randn(B, C, device="cuda") * 3).to(dtype).requires_grad_(True)
    target = torch.randint(0, C, (B,), device="cuda")
    loss = nat().cross_entropy(logits, target)
    loss.backward()
    ref_l = logits.detach().float().requires_grad_(True)
    ref = F.cross_entropy(ref_l, target)
    ref.backward()
    assert abs(loss.item() - ref.item()) < tol * max(1.0, abs(ref.item()))
    assert torch.allclose(logits.grad.float(), ref_l.grad, atol=tol / B, rtol=tol)


def test_count_correct():
    logits = torch.randn(257, 1000, device="cuda")
    target = logits.argmax(1)
    target[::4] = (target[::4] + 1) % 1000
    assert int(nat().count_correct(logits, target)) == int((logits.argmax(1) == target).sum())
    assert int(nat().count_correct(logits.bfloat16(), target)) == int((logits.bfloat16().float().argmax(1) == target).sum())


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-4), (torch.bfloat16, 3e-2)])
@pytest.mark.parametrize("C,relu,res", [(64, True, False), (128, True, True), (288, False, False), (80, True, True)])
def test_batchnorm_train_fwd_bwd(dtype, tol, C, relu, res):
    _bn_case(dtype, tol, C, relu, res, 8, 14, 14)


@pytest.mark.parametrize("C,relu,res,N,H,W", [(64, True, False, 35, 31, 31), (128, True, True, 16, 33, 33),
                                              (256, False, False, 16, 24, 24), (512, True, True, 9, 23, 23),
                                              (64, False, True, 43, 28, 28), (2048, True, False, 5, 15, 15)])
def test_batchnorm_streaming_kernels(C, relu, res, N, H, W):
    """bf16 tensors >= 4 MB take the bulk-copy streaming kernels (bn_stream.cu): ragged last chunks, every mask mode."""
    assert N * H * W * C * 2 >= 4 << 20
    _bn_case(torch.bfloat16, 3e-2, C, relu, res, N, H, W)


def _bn_case(dtype, tol, C, relu, res, N, H, W):
    torch.manual_seed(1)
    x = cl((torch.randn(N, C, H, W, device="cuda") * 2 + 0.5).to(dtype)).requires_grad_(True)
    r = cl(torch.randn(N, C, H, W, device="cuda").to(dtype)).requires_grad_(True) if res else None
    bn = nn.BatchNorm2d(C).cuda()
    with torch.no_grad():
        bn.weight.uniform_(0.5, 1.5)
        bn.bias.uniform_(-0.5, 0.5)
    bn_ref = copy.deepcopy(bn)
    out = nat().batch_norm_act(x, bn, r, relu)
    g = cl(torch.randn_like(out))
    out.backward(g)
    xr = x.detach().float().requires_grad_(True)
    rr = r.detach().float().requires_grad_(True) if res else None
    y = bn_ref(xr)
    if res:
        y = y + rr
    if relu:
        y = F.relu(y)
    y.backward(g.float())
    assert torch.allclose(out.float(), y, atol=tol * 4, rtol=tol)
    assert torch.allclose(x.grad.float(), xr.grad, atol=tol * 4, rtol=tol * 4)
    if res:
        assert torch.allclose(r.grad.float(), rr.grad, atol=tol, rtol=tol)
    assert torch.allclose(bn.weight.grad, bn_ref.weight.grad, atol=tol * 30, rtol=tol * 4)
    assert torch.allclose(bn.bias.grad, bn_ref.bias.grad, atol=tol * 30, rtol=tol * 4)
    assert torch.allclose(bn.running_mean, bn_ref.running_mean, atol=tol, rtol=tol)
    assert torch.allclose(bn.running_var, bn_ref.running_var, atol=tol, rtol=tol)
    assert int(bn.num_batches_tracked) == 1


def test_conv_bias_gradient_single_pass():
    """dL/db of a biased convolution (VGG / AlexNet / SqueezeNet) comes from one column-sum pass over dy."""
    torch.manual_seed(4)
    conv = nn.Conv2d(64, 128, 3, padding=1).cuda()
    x = cl(torch.randn(4, 64, 12, 12, device="cuda").bfloat16()).requires_grad_(True)
    y = nat().conv2d(x, conv.weight, conv.bias, (1, 1), (1, 1))
    g = cl(torch.randn_like(y))
    y.backward(g)
    xr = x.detach().float().requires_grad_(True)
    ref = copy.deepcopy(conv)
    ref.weight.grad = ref.bias.grad = None
    with torch.no_grad():
        ref.weight.copy_(conv.weight.bfloat16().float())
    F.conv2d(xr, ref.weight, ref.bias, 1, 1).backward(g.float())
    assert torch.allclose(conv.bias.grad, ref.bias.grad, atol=1e-3, rtol=1e-4)
    assert torch.allclose(conv.weight.grad, ref.weight.grad, atol=2e-2, rtol=2e-2)
    assert torch.allclose(x.grad.float(), xr.grad, atol=5e-2, rtol=5e-2)


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_channel_concat_and_slice_gather(dtype):
    """cat_channels writes each input into its NHWC channel slice; its backward hands out slice views that the
    consumers gather with the same row-strided copy kernel."""
    torch.manual_seed(6)
    n = nat()
    xs = [cl(torch.randn(3, c, 9, 7, device="cuda").to(dtype)).requires_grad_(True) for c in (64, 8, 96)]
    y = n.cat_channels(xs)
    assert y.is_contiguous(memory_format=torch.channels_last)
    assert torch.equal(y, torch.cat([x.detach() for x in xs], 1))
    z = n.relu(y)                                   # a consumer whose backward receives the full gradient
    g = cl(torch.randn_like(z))
    z.backward(g)
    off = 0
    for x in xs:
        c = x.shape[1]
        want = (g * (y > 0))[:, off:off + c]
        assert torch.equal(x.grad, want)
        off += c
    # a channel-slice view gathers into a dense NHWC tensor without the library's generic strided copy
    big = cl(torch.randn(2, 160, 5, 6, device="cuda").to(dtype))
    sl = big[:, 32:96]
    from mpi_pytorch_b200.ops import cuda_impl
    dense = cuda_impl._nhwc(sl)
    assert dense.is_contiguous(memory_format=torch.channels_last) and torch.equal(dense, sl)


def test_batchnorm_eval():
    C = 64
    x = cl(torch.randn(4, C, 7, 7, device="cuda"))
    bn = nn.BatchNorm2d(C).cuda().eval()
    with torch.no_grad():
        bn.running_mean.normal_()
        bn.running_var.uniform_(0.5, 2)
        out = nat().batch_norm_act(x, bn, None, True)
        assert torch.allclose(out, F.relu(bn(x)), atol=1e-5)


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-6), (torch.bfloat16, 1e-2)])
@pytest.mark.parametrize("k,s,p,ceil", [(3, 2, 1, False), (3, 2, 0, True), (2, 2, 0, False), (3, 2, 0, False)])
def test_maxpool(dtype, tol, k, s, p, ceil):
    torch.manual_seed(2)
    x = cl(F.relu(torch.randn(4, 64, 27, 27, device="cuda")).to(dtype)).requires_grad_(True)   # relu -> many ties
    y = nat().max_pool2d(x, k, s, p, ceil)
    g = cl(torch.randn_like(y))
    y.backward(g)
    xr = x.detach().float().requires_grad_(True)
    yr = F.max_pool2d(xr, k, s, p, 1, ceil)
    yr.backward(g.float())
    assert y.shape == yr.shape
    assert torch.allclose(y.float(), yr, atol=tol)
    assert torch.allclose(x.grad.float(), xr.grad, atol=tol * 4)


@pytest.mark.parametrize("dtype,tol", [(torch.float32, 1e-6), (torch.bfloat16, 1e-2)])
@pytest.mark.parametrize("hw", [(28, 28), (12, 30)])
def test_maxpool_stem_quad_backward(dtype, tol, hw):
    """3x3/s2/p1 on even H, W takes the one-thread-per-2x2-quad backward kernel (stem pool of ResNet/DenseNet)."""
    torch.manual_seed(5)
    x = cl(F.relu(torch.randn(3, 64, *hw, device="cuda")).to(dtype)).requires_grad_(True)
    y = nat().max_pool2d(x, 3, 2, 1, False)
    g = cl(torch.randn_like(y))
    y.backward(g)
    xr = x.detach().float().requires_grad_(True)
    yr = F.max_pool2d(xr, 3, 2, 1)
    yr.backward(g.float())
    assert torch.allclose(y.float(), yr, atol=tol)
    assert torch.allclose(x.grad.float(), xr.grad, atol=tol * 4)


@pytest.mark.parametrize("k,s,p,ceil,cip", [(2, 2, 0, False, True), (3, 1, 1, False, True), (5, 3, 0, False, True),
                                            (3, 2, 1, True, False)])
def test_avgpool(k, s, p, ceil, cip):
    x = cl(torch.randn(3, 32, 17, 17, device="cuda")).requires_grad_(True)
    y = nat().avg_pool2d(x, k, s, p, ceil, cip)
    g = cl(torch.randn_like(y))
    y.backward(g)
    xr = x.detach().cpu().contiguous().double().requires_grad_(True)      # CPU fp64 oracle
    yr = F.avg_pool2d(xr, k, s, p, ceil, cip)
    yr.backward(g.cpu().contiguous().double())
    assert torch.allclose(y.cpu().double(), yr, atol=1e-5)
    assert torch.allclose(x.grad.cpu().double(), xr.grad, atol=1e-5)


@pytest.mark.parametrize("out", [(1, 1), (6, 6), (7, 7)])
def test_adaptive_avgpool(out):
    x = cl(torch.randn(3, 64, 13, 13, device="cuda")).requires_grad_(True)
    y = nat().adaptive_avg_pool2d(x, out)
    g = cl(torch.randn_like(y))
    y.backward(g)
    xr = x.detach().clone().requires_grad_(True)
    yr = F.adaptive_avg_pool2d(xr, out)
    yr.backward(g)
    assert torch.allclose(y, yr, atol=1e-5) and torch.allclose(x.grad, xr.grad, atol=1e-5)


def test_relu_and_dropout():
    x = torch.randn(4, 64, 9, 9, device="cuda").requires_grad_(True)
    y = nat().relu(x)
    y.sum().backward()
    assert torch.equal(y, F.relu(x)) and torch.equal(x.grad, (x > 0).float())
    z = torch.ones(1 << 16, device="cuda", requires_grad=True)
    d = nat().dropout(z, 0.5)
    keep = (d != 0).float().mean().item()
    assert 0.47 < keep < 0.53 and torch.allclose(d[d != 0], torch.tensor(2.0, device="cuda"))
    d.sum().backward()
    assert torch.equal(z.grad, d.detach())


def test_dropout_in_cuda_graph_draws_a_new_mask_per_replay():
    """The dropout seed is mixed with a device-resident step counter, so a captured step is replayable."""
    n = nat()
    x = torch.ones(64 * 1024, device="cuda", requires_grad=True)
    n.dropout_epoch(x.device)                       # the counter must exist before capture
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):                   # warm-up outside the graph
        torch.autograd.grad(n.dropout(x, 0.5).sum(), x)
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        n.advance_dropout_epoch(x.device)
        y = n.dropout(x, 0.5)
        (gx,) = torch.autograd.grad(y.sum(), x)
    masks = []
    for _ in range(3):
        g.replay()
        torch.cuda.synchronize()
        assert torch.equal(y != 0, gx != 0)         # forward and backward of one replay agree
        assert 0.45 < float((y != 0).float().mean()) < 0.55
        assert torch.all(y[y != 0] == 2.0)
        masks.append((y != 0).clone())
    assert not torch.equal(masks[0], masks[1]) and not torch.equal(masks[1], masks[2])


def test_adam_and_sgd_flat():
    from mpi_pytorch_b200.ops import reference as ref
    torch.manual_seed(3)
    n = 100003
    p = torch.randn(n, device="cuda"); g = torch.randn(n, device="cuda")
    m = torch.zeros(n, device="cuda"); v = torch.zeros(n, device="cuda")
    p2, m2, v2 = p.clone(), m.clone(), v.clone()
    sh = torch.empty(n, dtype=torch.bfloat16, device="cuda")
    for step in (1, 2, 3):
        nat().adam_flat(p, g, m, v, sh, step, 4e-4, 0.9, 0.999, 1e-8, 0.0)
        ref.adam_step(p2, g, m2, v2, step, 4e-4)
    assert torch.allclose(p, p2, atol=1e-6) and torch.allclose(m, m2, atol=1e-6) and torch.allclose(v, v2, atol=1e-8, rtol=1e-4)
    assert torch.equal(sh, p.bfloat16())
    q = torch.randn(n, device="cuda"); q2 = q.clone(); b = torch.zeros(n, device="cuda"); b2 = b.clone()
    for step in (1, 2):
        nat().sgd_flat(q, g, b, None, step, 0.1, 0.9, 0.0)
        ref.sgd_step(q2, g, b2, step, 0.1, 0.9)
    assert torch.allclose(q, q2, atol=1e-6)


@pytest.mark.parametrize("mode", ["bilinear", "bicubic"])
def test_resize_normalize(mode):
    from mpi_pytorch_b200.ops import reference as ref
    torch.manual_seed(4)
    img = torch.randint(0, 256, (2, 250, 170, 3), dtype=torch.uint8, device="cuda")
    out = nat().resize_normalize(img, 128, 128, mode, torch.float32, False)
    want = ref.resize_normalize(img.cpu(), 128, 128, mode).cuda()
    # one uint8 quantisation step of slack for the bicubic/PIL path (1/255/0.225 = 0.0175)
    assert (out - want).abs().max().item() < (0.02 if mode == "bicubic" else 2e-3)
    u8 = nat().resize_u8(img, 64, 64, mode)
    want8 = ref.resize_u8(img.cpu(), 64, 64, mode).cuda()
    assert (u8.int() - want8.int()).abs().max().item() <= 1
    nz = nat().normalize_u8(u8, torch.float32)
    assert torch.allclose(nz, ref.normalize_u8(u8.cpu()).cuda(), atol=1e-5)


def test_synth_matches_torch_formula():
    from mpi_pytorch_b200.data.synthetic import synth_images_torch
    idx = torch.tensor([0, 5, 123456789, 7], device="cuda")
    a = nat().synth_batch(3, idx, 3, 16, 20, torch.float32, False)
    b = synth_images_torch(3, idx.cpu(), 3, 16, 20).cuda()
    assert torch.allclose(a, b, atol=1e-6)
    c = nat().synth_batch(3, idx, 3, 16, 20, torch.float32, True)
    assert c.is_contiguous(memory_format=torch.channels_last) and torch.allclose(c, b, atol=1e-6)


def test_peer_collectives_world1_fused_adam():
    """W=1 degenerate run of the fused all-reduce+Adam kernel (sync logic is covered at N=2 by scripts/test_peer.py)."""
    from mpi_pytorch_b200.ops import reference as ref
    from mpi_pytorch_b200.parallel import bootstrap
    from mpi_pytorch_b200.parallel.collectives import Hyper
    from mpi_pytorch_b200.parallel.peer import PeerCollectives
    w = bootstrap.init_world("cuda")
    pc = PeerCollectives(w)
    for n, algo in ((4096, "oneshot"), (1 << 20, "twoshot")):
        g = pc.alloc(n, torch.float32); p = pc.alloc(n, torch.float32); sh = pc.alloc(n, torch.bfloat16)
        g.normal_(); p.normal_()
        m = torch.zeros(n, device="cuda"); v = torch.zeros(n, device="cuda")
        p2, m2, v2 = p.clone(), m.clone(), v.clone()
        h = Hyper("adam", lr=1e-3)
        for step in (1, 2):
            pc.all_reduce_update_(g, p, {"exp_avg": m, "exp_avg_sq": v}, sh, h, step, algo)
            ref.adam_step(p2, g, m2, v2, step, 1e-3)
        torch.cuda.synchronize()
        assert torch.allclose(p, p2, atol=1e-6) and torch.equal(sh, p.bfloat16())
        g0 = g.clone()
        pc.all_reduce_avg_(g, algo)
        torch.cuda.synchronize()
        assert torch.equal(g, g0)


def test_trainer_native_step_matches_torch_path():
    """One ResNet-18 step on the native path vs the torch/cuDNN path, fp32 mode."""
    from mpi_pytorch_b200 import ops
    from mpi_pytorch_b200.engine.trainer import Trainer
    from mpi_pytorch_b200.parallel import bootstrap
    from mpi_pytorch_b200.utils.config import Settings
    w = bootstrap.init_world("cuda")
    s = Settings(MODEL_NAME="resnet18", NUM_CLASSES=16, BATCH_SIZE=8, HEIGHT=64, WIDTH=64, USE_PRETRAINED=False,
                 DTYPE="fp32", LR=1e-3)
    losses = {}
    for native in (True, False):
        ops.set_native(native)
        try:
            tr = Trainer(s, w)
            shard = tr.make_shard(total=16, shuffle=False)
            losses[native] = [float(tr.train_step(x, y)) for x, y in shard]
            tr.close()
        finally:
            ops.set_native(True)
    assert abs(losses[True][0] - losses[False][0]) < 1e-3
    assert abs(losses[True][1] - losses[False][1]) < 5e-2
