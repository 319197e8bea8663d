"""CUDA execution path: ctypes bindings to ``libmpb200.so`` + autograd Functions.

Layout contract: 4-D activations are channels-last (physical NHWC), dtype bf16 (fp32 is supported
by every kernel too and is what the numerics tests use as "fp32 mode").  Parameters are fp32 masters;
when the DP engine attached a bf16 shadow (``p._mpb_shadow``, refreshed by the fused all-reduce+Adam
kernel) the compute kernels read that instead of casting.

Importing this module on a machine without the built library raises: there is no silent fallback
(the driver checks that the in-tree ``.so`` was really loaded by the GPU tests).
"""
from __future__ import annotations

import ctypes
import itertools
import os
from typing import Optional, Tuple

import torch
import torch.nn.functional as F

from . import scratch as _scratch

from ..build import lib_path

_LIB = ctypes.CDLL(lib_path())
LAUNCHES = {"n": 0}            # kernels launched through this module (bench.py's gpu_launches)
FALLBACKS = {}                 # op -> count of calls that went to a library kernel (cuDNN/cuBLAS)

c_void_p, c_int, c_ll, c_float = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float


def _dt(t: torch.Tensor) -> int:
    if t.dtype == torch.bfloat16:
        return 1
    if t.dtype == torch.float32:
        return 0
    raise TypeError("mpb kernels take fp32 or bf16, got {}".format(t.dtype))


def _stream() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: Optional[torch.Tensor]):
    return c_void_p(0 if t is None else t.data_ptr())


def _check(rc: int, name: str) -> None:
    if rc != 0:
        raise RuntimeError("{} failed with code {}".format(name, rc))


def _call(name: str, *args, launches: int = 1) -> None:
    rc = getattr(_LIB, name)(*args)
    LAUNCHES["n"] += launches
    _check(rc, name)


def _note_fallback(op: str) -> None:
    FALLBACKS[op] = FALLBACKS.get(op, 0) + 1


def _row_view(t: torch.Tensor):
    """(rows, channels, channel pitch) if ``t`` (NCHW-shaped) is stored as NHWC rows with a channel pitch >= C -- i.e.
    a channel slice of a wider channels-last tensor (what the backward of a channel concat hands out)."""
    if t.dim() != 4:
        return None
    n, c, h, w = t.shape
    sn, sc, sh, sw = t.stride()
    if (c > 1 and sc != 1) or sw < c or (h > 1 and sh != w * sw) or (n > 1 and sn != h * w * sw):
        return None
    return n * h * w, c, sw


def _copy_rows(src_ptr: int, src_pitch: int, dst_ptr: int, dst_pitch: int, rows: int, row_bytes: int) -> None:
    _call("mpb_copy_rows", c_void_p(src_ptr), c_ll(src_pitch), c_void_p(dst_ptr), c_ll(dst_pitch), c_ll(rows),
          c_ll(row_bytes), c_void_p(_stream()))


def _nhwc(x: torch.Tensor) -> torch.Tensor:
    if x.dim() == 4 and not x.is_contiguous(memory_format=torch.channels_last):
        rv = _row_view(x) if x.is_cuda else None
        eb = x.element_size()
        if rv is not None and (rv[1] * eb) % 16 == 0 and (rv[2] * eb) % 16 == 0 and x.data_ptr() % 16 == 0:
            out = torch.empty(x.shape, dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
            _copy_rows(x.data_ptr(), rv[2] * eb, out.data_ptr(), rv[1] * eb, rv[0], rv[1] * eb)     # 16 B per thread
            return out
        return x.contiguous(memory_format=torch.channels_last)
    return x


def _pair(v) -> Tuple[int, int]:
    return (int(v[0]), int(v[1])) if isinstance(v, (tuple, list)) else (int(v), int(v))


def _f32(t: Optional[torch.Tensor]) -> Optional[torch.Tensor]:
    """Bias / affine vectors are read by the kernels as fp32: cast the rare all-bf16 module instead of misreading it."""
    if t is None:
        return None
    t = t.detach()
    return t if t.dtype == torch.float32 else t.float()


def prepare_inference(model: torch.nn.Module, dtype: torch.dtype = torch.bfloat16) -> torch.nn.Module:
    """Make a model ready for forward-only use on the native path WITHOUT the DP engine: parameters, BatchNorm
    statistics and biases stay fp32 (what the kernels read), and every GEMM-shaped weight gets the bf16 compute copy
    the engine would otherwise maintain, so no per-call cast happens (eval pipeline predictors,
    /root/reference/evaluation_pipeline.py:138-147)."""
    with torch.no_grad():
        for p in model.parameters():
            if p.dim() >= 2 and p.dtype == torch.float32 and dtype != torch.float32:
                if p.dim() == 4:
                    p._mpb_shadow = p.detach().to(dtype).contiguous(memory_format=torch.channels_last)
                else:
                    p._mpb_shadow = p.detach().to(dtype).contiguous()
        for m in model.modules():
            if isinstance(m, torch.nn.Conv2d):
                m._mpb_fold_ok = True      # weights are frozen from here on: conv_bn_act may fold BatchNorm into them
                m._mpb_fold = None
    return model.eval()


def _compute_weight(w: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    sh = getattr(w, "_mpb_shadow", None)
    if sh is not None and sh.dtype == dtype:
        return sh
    return w.detach().to(dtype) if w.dtype != dtype else w.detach()


# =============================================================================== convolution / linear
def _gemm():
    from . import tc
    return tc


class _ConvFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, stride, padding, dilation, groups, want_stats, relu=False):
        # the (non-differentiable) statistics output would otherwise get a materialised zeros gradient: one tiny
        # fill launch per convolution per step
        ctx.set_materialize_grads(False)
        x = _nhwc(x)
        w = _compute_weight(weight, x.dtype)
        ctx.conf = (stride, padding, dilation, groups, bias is not None, weight.dtype)
        ctx.weight_param = weight if getattr(weight, "_mpb_grad_ready", None) is not None else None
        tc = _gemm()
        stats = None
        ctx.xcol = None
        ctx.xs2d = None
        fused_relu = False
        if tc.stem_s2d_ok(x, w, stride, padding, dilation, groups):
            y, stats, ctx.xs2d = tc.conv_fprop_stem(x, w, _f32(bias), padding, want_stats)
        elif tc.fprop_ok(x, w, stride, padding, dilation, groups):
            # Conv2d -> ReLU pairs: the activation runs in the conv epilogue (statistics, when wanted, are of the
            # pre-activation values, so the two are never combined)
            y, stats = tc.conv_fprop(x, w, _f32(bias), stride, padding, want_stats, relu=relu and not want_stats)
            fused_relu = relu and not want_stats
        elif tc.im2col_ok(x, w, stride, padding, dilation, groups):
            y, stats, ctx.xcol = tc.conv_fprop_im2col(x, w, _f32(bias), stride, padding,
                                                      want_stats)
        else:
            _note_fallback("conv_fprop")
            b = None if bias is None else bias.detach().to(x.dtype)
            y = _nhwc(F.conv2d(x, w, b, stride, padding, dilation, groups))
            if want_stats:
                stats = bn_stats(y)
        if relu:
            if not fused_relu:
                _call("mpb_relu_fwd", _p(y), _p(y), c_ll(y.numel()), c_int(_dt(y)), c_void_p(_stream()))   # in place
            ctx.save_for_backward(x, w, y)   # y is the mask source of the ReLU backward
        else:
            ctx.save_for_backward(x, w)
        ctx.mark_non_differentiable(*([stats] if stats is not None else []))
        if want_stats:
            return y, stats
        return y

    @staticmethod
    def backward(ctx, gy, *unused):
        if gy is None:
            return (None,) * 9
        saved = ctx.saved_tensors
        x, w = saved[0], saved[1]
        stride, padding, dilation, groups, has_bias, wdtype = ctx.conf
        gy = _nhwc(gy)
        gb_fused = None
        if len(saved) > 2:
            # ReLU backward; for a biased conv the same pass also produces the bias gradient (column sums of dy)
            yo, k = saved[2], gy.shape[1]
            dy = torch.empty_like(gy)
            if has_bias and ctx.needs_input_grad[2] and k % 8 == 0 and k <= 2048:
                sums = _scratch.take(2 * k, gy.device)[0]
                _call("mpb_relu_bwd_colsum", _p(gy), _p(yo), _p(dy), _p(sums), c_ll(gy.numel() // k), c_int(k),
                      c_int(_dt(gy)), c_void_p(_stream()))
                gb_fused = sums[:k]
            else:
                _call("mpb_relu_bwd", _p(gy), _p(yo), _p(dy), c_ll(gy.numel()), c_int(_dt(gy)), c_void_p(_stream()))
            gy = dy
        need_x, need_w = ctx.needs_input_grad[0], ctx.needs_input_grad[1]
        tc = _gemm()
        gx = gw = gb = None
        args = (x, w, stride, padding, dilation, groups)
        if need_x and tc.dgrad_ok(*args):
            gx = tc.conv_dgrad(gy, w, x.shape, stride, padding)
            need_x = False
        if need_w and ctx.xs2d is not None:
            wp = ctx.weight_param
            if wp is not None and wp.grad is not None and wp.grad.dtype == torch.float32 \
                    and wp.grad.permute(0, 2, 3, 1).is_contiguous():
                tc.conv_wgrad_stem(gy, ctx.xs2d, x.shape, w.shape, padding, out=wp.grad)   # += the arena slot
                wp._mpb_grad_ready(wp)
            else:
                gw = tc.conv_wgrad_stem(gy, ctx.xs2d, x.shape, w.shape, padding).to(wdtype)
            need_w = False
        if need_w and ctx.xcol is not None:
            gw = tc.conv_wgrad_im2col(gy, ctx.xcol, w.shape).to(wdtype)
            need_w = False
        if need_w and tc.wgrad_ok(*args):
            wp = ctx.weight_param
            if wp is not None and wp.grad is not None and wp.grad.dtype == torch.float32 \
                    and wp.grad.permute(0, 2, 3, 1).is_contiguous():
                # accumulate straight into the parameter's slot of the (zeroed) gradient arena and tell the
                # DP engine the gradient is complete; autograd sees "no gradient" for the weight
                tc.conv_wgrad(gy, x, w.shape, stride, padding, out=wp.grad)
                wp._mpb_grad_ready(wp)
            else:
                gw = tc.conv_wgrad(gy, x, w.shape, stride, padding).to(wdtype)
            need_w = False
        if need_x or need_w:
            _note_fallback("conv_dgrad" if need_x else "conv_wgrad")
            if need_x and need_w:
                _note_fallback("conv_wgrad")
            fx, fw, _ = torch.ops.aten.convolution_backward(
                gy, x, w, None, list(stride), list(padding), list(dilation), False, [0, 0], groups,
                [need_x, need_w, False])
            if need_x:
                gx = fx
            if need_w:
                gw = fw.to(wdtype)
        if has_bias and ctx.needs_input_grad[2]:
            k = gy.shape[1]
            if gb_fused is not None:
                gb = gb_fused
            elif k % 8 == 0 and k <= 2048 and gy.dtype in (torch.bfloat16, torch.float32):
                gb = bn_stats(gy)[:k]             # one pass over dy in its own dtype (no fp32 copy + library reduction)
            else:
                gb = gy.float().sum(dim=(0, 2, 3))
        return gx, gw, gb, None, None, None, None, None, None


def conv2d(x, w, b, stride, padding, dilation=(1, 1), groups=1):
    if x.dtype not in (torch.bfloat16, torch.float32):
        return F.conv2d(x, w, b, stride, padding, dilation, groups)
    return _ConvFn.apply(x, w, b, _pair(stride), _pair(padding), _pair(dilation), int(groups), False)


def conv_relu(x, conv):
    """Conv2d -> ReLU as one unit: activation in the conv epilogue, mask + bias gradient in one backward pass."""
    if x.dtype != torch.bfloat16 or x.dim() != 4:
        return relu(conv2d(x, conv.weight, conv.bias, conv.stride, conv.padding, conv.dilation, conv.groups))
    return _ConvFn.apply(x, conv.weight, conv.bias, _pair(conv.stride), _pair(conv.padding), _pair(conv.dilation),
                         int(conv.groups), False, True)


def _arena_slot_ok(p) -> bool:
    """True when ``p`` is an engine-managed parameter whose gradient is a contiguous fp32 view of the (zeroed) gradient
    arena: the dW kernel may then add into it directly and report completion, instead of returning a tensor that
    autograd would add with one more pass."""
    return (p is not None and p.grad is not None and p.grad.dtype == torch.float32 and p.grad.dim() == 2
            and p.grad.is_contiguous())


class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias):
        w = _compute_weight(weight, x.dtype)
        ctx.save_for_backward(x, w)
        ctx.conf = (bias is not None, weight.dtype)
        ctx.weight_param = weight if getattr(weight, "_mpb_grad_ready", None) is not None else None
        tc = _gemm()
        if tc.gemm_supported(x, w):
            return tc.linear_fwd(x, w, _f32(bias))
        _note_fallback("linear_fwd")
        return F.linear(x, w, None if bias is None else bias.detach().to(x.dtype))

    @staticmethod
    def backward(ctx, gy):
        x, w = ctx.saved_tensors
        has_bias, wdtype = ctx.conf
        gy = gy.contiguous()
        tc = _gemm()
        gx = gw = gb = None
        if tc.gemm_supported(x, w):
            if ctx.needs_input_grad[0]:
                gx = tc.linear_dx(gy, w)
            if ctx.needs_input_grad[1]:
                wp = ctx.weight_param
                if _arena_slot_ok(wp):
                    tc.linear_dw(gy, x, out=wp.grad)            # += the weight's slot of the gradient arena
                    wp._mpb_grad_ready(wp)
                else:
                    gw = tc.linear_dw(gy, x).to(wdtype)
        else:
            _note_fallback("linear_bwd")
            if ctx.needs_input_grad[0]:
                gx = gy @ w
            if ctx.needs_input_grad[1]:
                gw = (gy.t() @ x).to(wdtype)
        if has_bias and ctx.needs_input_grad[2]:
            k = gy.shape[1]
            if k % 8 == 0 and k <= 2048 and gy.dtype in (torch.bfloat16, torch.float32):
                gb = bn_stats(gy.view(gy.shape[0], k, 1, 1))[:k]     # the column-sum kernel conv biases use
            else:
                gb = gy.float().sum(dim=0)
        return gx, gw, gb


class _LinearPadFn(torch.autograd.Function):
    """Linear layer whose output width is not a multiple of 8 -- the reference's default head has 64,500 classes
    (utils.py:39) -- on the tensor-core GEMM: TMA needs 16-byte row strides, so the layer is computed with the class
    dimension zero-padded to the next multiple of 8 (compute copy of W, logits, their gradient) and sliced back."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        n, k = weight.shape
        npad = (n + 7) // 8 * 8
        w = _compute_weight(weight, x.dtype)
        wp = torch.zeros((npad, k), dtype=x.dtype, device=x.device)
        wp[:n].copy_(w)
        bp = None
        if bias is not None:
            bp = torch.zeros(npad, dtype=torch.float32, device=x.device)
            bp[:n].copy_(bias.detach())
        ctx.save_for_backward(x, wp)
        ctx.conf = (bias is not None, weight.dtype, n)
        ctx.weight_param = weight if getattr(weight, "_mpb_grad_ready", None) is not None else None
        return _gemm().linear_fwd(x, wp, bp)[:, :n]

    @staticmethod
    def backward(ctx, gy):
        x, wp = ctx.saved_tensors
        has_bias, wdtype, n = ctx.conf
        tc = _gemm()
        gp = torch.zeros((gy.shape[0], wp.shape[0]), dtype=x.dtype, device=x.device)
        gp[:, :n].copy_(gy)
        gx = tc.linear_dx(gp, wp) if ctx.needs_input_grad[0] else None
        gw = None
        if ctx.needs_input_grad[1]:
            wparam = ctx.weight_param
            if _arena_slot_ok(wparam):
                tc.linear_dw(gp, x, out=wparam.grad, rows=n)      # only the n real rows are written
                wparam._mpb_grad_ready(wparam)
            else:
                gw = tc.linear_dw(gp, x)[:n].to(wdtype)
        gb = gy.float().sum(dim=0) if (has_bias and ctx.needs_input_grad[2]) else None
        return gx, gw, gb


def linear(x, w, b):
    if x.dtype not in (torch.bfloat16, torch.float32) or x.dim() != 2:
        return F.linear(x, w.to(x.dtype), None if b is None else b.to(x.dtype))
    if (w.shape[0] % 8 != 0 and w.shape[1] % 8 == 0 and x.shape[0] % 8 == 0 and x.dtype == torch.bfloat16
            and _gemm().ENABLED):
        return _LinearPadFn.apply(x.contiguous(), w, b)
    return _LinearFn.apply(x.contiguous(), w, b)


# =============================================================================== batch norm (+add +relu)
class _BNActFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, residual, run_mean, run_var, eps, momentum, relu, training, sums, nbt=None):
        x = _nhwc(x)
        n, c, h, w = x.shape
        m = n * h * w
        res = _nhwc(residual) if residual is not None else None
        out = torch.empty_like(x)
        dev = x.device
        save_mean = torch.empty(c, dtype=torch.float32, device=dev)
        save_invstd = torch.empty(c, dtype=torch.float32, device=dev)
        if training and sums is None:
            sums = torch.empty(2 * c, dtype=torch.float32, device=dev)
            _call("mpb_bn_stats", _p(x), _p(sums), c_ll(m), c_int(c), c_int(_dt(x)), c_void_p(_stream()), launches=2)
        _call("mpb_bn_fwd_apply_ex", _p(x), _p(res), _p(out), _p(sums), _p(gamma), _p(beta), _p(run_mean), _p(run_var),
              _p(save_mean), _p(save_invstd), _p(nbt), c_ll(m), c_int(c), c_float(eps), c_float(momentum),
              c_int(int(relu)), c_int(int(training)), c_int(_dt(x)), c_void_p(_stream()))
        # affine gradients go straight into the DP engine's gradient arena when the parameters live there
        ctx.direct = None
        if (gamma is not None and beta is not None and getattr(gamma, "_mpb_grad_ready", None) is not None
                and getattr(beta, "_mpb_grad_ready", None) is not None):
            ctx.direct = (gamma, beta)
        # BN -> ReLU without a residual: backward recomputes the mask from x, so `out` need not be kept for it
        recompute = bool(relu) and residual is None and training and beta is not None
        ctx.save_for_backward(x, None if recompute else out, gamma, beta if recompute else None, save_mean, save_invstd)
        ctx.conf = (relu, residual is not None, training, c, m)
        return out

    @staticmethod
    def backward(ctx, gz):
        x, out, gamma, beta, save_mean, save_invstd = ctx.saved_tensors
        relu, has_res, training, c, m = ctx.conf
        if not training:
            raise RuntimeError("backward through eval-mode BatchNorm is not supported on the native path")
        gz = _nhwc(gz)
        if gz.dtype != x.dtype:
            gz = gz.to(x.dtype)
        dx = torch.empty_like(x)
        dsum, pooled = _scratch.take(2 * c, x.device)
        direct = None
        if ctx.direct is not None:
            gp, bp = ctx.direct
            if all(t.grad is not None and t.grad.dtype == torch.float32 and t.grad.is_contiguous() for t in (gp, bp)):
                direct = (gp, bp)
        if direct is not None:
            dgamma, dbeta, acc = direct[0].grad, direct[1].grad, 1
        else:
            dgamma = torch.empty(c, dtype=torch.float32, device=x.device)
            dbeta = torch.empty(c, dtype=torch.float32, device=x.device)
            acc = 0
        acc |= 2          # dsum is zero already (pool slice or fresh torch.zeros)
        dres = None
        if out is None:                                   # recompute-mask variant: reads dz and x only
            _call("mpb_bn_relu_bwd_recompute_ex", _p(gz), _p(x), _p(gamma), _p(beta), _p(save_mean), _p(save_invstd),
                  _p(dsum), _p(dx), _p(dgamma), _p(dbeta), c_ll(m), c_int(c), c_int(acc), c_int(_dt(x)),
                  c_void_p(_stream()), launches=2)
        else:
            dres = torch.empty_like(x) if has_res else None
            _call("mpb_bn_bwd_ex", _p(gz), _p(out), _p(x), _p(gamma), _p(save_mean), _p(save_invstd), _p(dsum), _p(dx),
                  _p(dres), _p(dgamma), _p(dbeta), c_ll(m), c_int(c), c_int(int(relu)), c_int(acc), c_int(_dt(x)),
                  c_void_p(_stream()), launches=2)
            if has_res and not relu:
                dres = gz
        if direct is not None:
            for t in direct:
                t._mpb_grad_ready(t)
            dgamma = dbeta = None
        return dx, dgamma, dbeta, dres, None, None, None, None, None, None, None, None


def bn_stats(y: torch.Tensor) -> torch.Tensor:
    """Per-channel (sum, sum of squares) of an NHWC tensor -> fp32 [2C] (what the conv epilogue fuses)."""
    n, c, h, w = y.shape
    sums = torch.empty(2 * c, dtype=torch.float32, device=y.device)
    _call("mpb_bn_stats", _p(y), _p(sums), c_ll(n * h * w), c_int(c), c_int(_dt(y)), c_void_p(_stream()), launches=2)
    return sums


def _bn_ok(x, bn) -> bool:
    c = x.shape[1]
    return (x.dim() == 4 and c % 8 == 0 and c <= 2048 and x.dtype in (torch.bfloat16, torch.float32)
            and bn.affine and bn.track_running_stats and bn.weight.dtype == torch.float32
            and bn.running_mean.dtype == torch.float32)


def batch_norm_act(x, bn, residual=None, relu=False, sums=None):
    if not _bn_ok(x, bn) or (not bn.training and torch.is_grad_enabled() and x.requires_grad):
        from . import reference as ref
        _note_fallback("batch_norm")
        return ref.batch_norm_act(x, bn, residual, relu)
    momentum = 0.1 if bn.momentum is None else float(bn.momentum)
    nbt = bn.num_batches_tracked
    if nbt is not None and not (nbt.is_cuda and nbt.dtype == torch.int64):
        if bn.training:
            nbt.add_(1)
        nbt = None
    return _BNActFn.apply(x, bn.weight, bn.bias, residual, bn.running_mean, bn.running_var, float(bn.eps), momentum,
                          bool(relu), bool(bn.training), sums, nbt)


FOLDED = {"n": 0}              # conv+BN(+ReLU) calls served by ONE kernel through folded weights (inference)


def _folded(conv, bn, dtype):
    """Eval-mode BatchNorm folded into the convolution: w' = w * g / sqrt(var + eps) per output channel,
    b' = beta - mean * g / sqrt(var + eps) (+ the conv bias scaled alike).  Built once per (conv, bn) pair of a
    model that ``prepare_inference`` froze; the fprop kernel then applies b' and the ReLU in its epilogue."""
    fold = getattr(conv, "_mpb_fold", None)
    if fold is None or fold[0] is not bn:
        with torch.no_grad():
            scale = bn.weight.float() * torch.rsqrt(bn.running_var.float() + bn.eps)
            w = (conv.weight.float() * scale.view(-1, 1, 1, 1)).to(dtype).contiguous(memory_format=torch.channels_last)
            b = bn.bias.float() - bn.running_mean.float() * scale
            if conv.bias is not None:
                b = b + conv.bias.float() * scale
        fold = conv._mpb_fold = (bn, w, b.contiguous())
    return fold[1], fold[2]


def conv_bn_act(x, conv, bn, relu=True, residual=None):
    if x.dtype not in (torch.bfloat16, torch.float32):
        x = x.to(torch.bfloat16)
    if (not bn.training and residual is None and getattr(conv, "_mpb_fold_ok", False) and not torch.is_grad_enabled()
            and x.dtype == torch.bfloat16 and bn.track_running_stats and bn.affine):
        tc = _gemm()
        x = _nhwc(x)
        stride, padding, dil = _pair(conv.stride), _pair(conv.padding), _pair(conv.dilation)
        w, b = _folded(conv, bn, x.dtype)
        if tc.fprop_ok(x, w, stride, padding, dil, int(conv.groups)) and \
                not tc.stem_s2d_ok(x, w, stride, padding, dil, int(conv.groups)):
            FOLDED["n"] += 1
            return tc.conv_fprop(x, w, b, stride, padding, False, relu=relu)[0]
    want_stats = bn.training and _bn_ok_channels(conv.out_channels)
    res = _ConvFn.apply(x, conv.weight, conv.bias, _pair(conv.stride), _pair(conv.padding), _pair(conv.dilation),
                        int(conv.groups), want_stats)
    y, sums = res if want_stats else (res, None)
    return batch_norm_act(y, bn, residual, relu, sums)


def _bn_ok_channels(c: int) -> bool:
    return c % 8 == 0 and c <= 2048


# =============================================================================== channel concat
class _CatFn(torch.autograd.Function):
    """Channel concat of NHWC tensors: each input is written into its channel slice with the row-strided copy kernel;
    backward gathers each input's channel slice of the gradient into a dense tensor with the same kernel, so whatever
    consumes it next -- one of our kernels or autograd's own accumulation (DenseNet: O(L^2) adds) -- runs vectorised."""

    @staticmethod
    def forward(ctx, *tensors):
        ts = [_nhwc(t) for t in tensors]
        n, _, h, w = ts[0].shape
        cs = [int(t.shape[1]) for t in ts]
        ct, eb = sum(cs), ts[0].element_size()
        out = torch.empty((n, ct, h, w), dtype=ts[0].dtype, device=ts[0].device, memory_format=torch.channels_last)
        off = 0
        for t, c in zip(ts, cs):
            _copy_rows(t.data_ptr(), c * eb, out.data_ptr() + off * eb, ct * eb, n * h * w, c * eb)
            off += c
        ctx.cs = cs
        return out

    @staticmethod
    def backward(ctx, gy):
        outs, off = [], 0
        for i, c in enumerate(ctx.cs):
            outs.append(_nhwc(gy[:, off:off + c]) if ctx.needs_input_grad[i] else None)
            off += c
        return tuple(outs)


def cat_channels(tensors):
    t0 = tensors[0]
    ok = all(t.dim() == 4 and t.is_cuda and t.dtype == t0.dtype and t.shape[0] == t0.shape[0] and t.shape[2:] == t0.shape[2:]
             and (t.shape[1] * t.element_size()) % 16 == 0 for t in tensors)
    if not ok or t0.dtype not in (torch.bfloat16, torch.float32, torch.float16):
        return torch.cat(list(tensors), dim=1)
    return _CatFn.apply(*tensors)


# =============================================================================== elementwise
class _ReluFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        y = torch.empty_like(x)
        _call("mpb_relu_fwd", _p(x), _p(y), c_ll(x.numel()), c_int(_dt(x)), c_void_p(_stream()))
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, gy):
        (y,) = ctx.saved_tensors
        gy = _nhwc(gy) if (y.dim() == 4 and y.is_contiguous(memory_format=torch.channels_last)) else gy.contiguous()
        dx = torch.empty_like(y)
        _call("mpb_relu_bwd", _p(gy), _p(y), _p(dx), c_ll(y.numel()), c_int(_dt(y)), c_void_p(_stream()))
        return dx


def _dense(x: torch.Tensor) -> bool:
    return x.is_contiguous() or (x.dim() == 4 and x.is_contiguous(memory_format=torch.channels_last))


def relu(x, inplace=False):
    if x.numel() % 8 != 0 or x.dtype not in (torch.bfloat16, torch.float32) or not _dense(x):
        return F.relu(x, inplace=inplace)
    return _ReluFn.apply(x)


_DROP_SEED = itertools.count(0x5EED)


_DROP_EPOCH = {}      # device -> int64[1] step counter mixed into every dropout seed ON THE DEVICE


def dropout_epoch(device) -> torch.Tensor:
    dev = torch.device(device)
    if dev not in _DROP_EPOCH:
        _DROP_EPOCH[dev] = torch.zeros(1, dtype=torch.int64, device=dev)
    return _DROP_EPOCH[dev]


def advance_dropout_epoch(device) -> None:
    """One launch per training step (captured with it): replays of a CUDA graph then draw fresh dropout masks even
    though the host-side seeds were frozen at capture time."""
    dropout_epoch(device).add_(1)


class _DropoutFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, p, seed):
        y = torch.empty_like(x)
        ep = dropout_epoch(x.device)
        _call("mpb_dropout_ex", _p(x), _p(y), c_ll(x.numel()), ctypes.c_uint(seed), _p(ep), c_float(p), c_int(_dt(x)),
              c_void_p(_stream()))
        ctx.conf = (p, seed)
        return y

    @staticmethod
    def backward(ctx, gy):
        p, seed = ctx.conf
        gy = gy.contiguous()
        dx = torch.empty_like(gy)
        ep = dropout_epoch(gy.device)
        _call("mpb_dropout_ex", _p(gy), _p(dx), c_ll(gy.numel()), ctypes.c_uint(seed), _p(ep), c_float(p),
              c_int(_dt(gy)), c_void_p(_stream()))
        return dx, None, None


def dropout(x, p):
    if x.numel() % 8 != 0 or x.dtype not in (torch.bfloat16, torch.float32) or not x.is_contiguous():
        return F.dropout(x, p, True)
    return _DropoutFn.apply(x, float(p), next(_DROP_SEED) & 0xFFFFFFFF)


# =============================================================================== pooling
def _out_size(i: int, k: int, s: int, p: int, ceil_mode: bool) -> int:
    if ceil_mode:
        o = -(-(i + 2 * p - k) // s) + 1
        if (o - 1) * s >= i + p:
            o -= 1
    else:
        o = (i + 2 * p - k) // s + 1
    return o


class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, s, p, ceil_mode):
        x = _nhwc(x)
        n, c, h, w = x.shape
        ho, wo = _out_size(h, k[0], s[0], p[0], ceil_mode), _out_size(w, k[1], s[1], p[1], ceil_mode)
        y = torch.empty((n, c, ho, wo), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        idx = torch.empty((n, ho, wo, c), dtype=torch.uint8, device=x.device)
        _call("mpb_maxpool_fwd", _p(x), _p(y), _p(idx), c_int(n), c_int(h), c_int(w), c_int(c), c_int(ho), c_int(wo),
              c_int(k[0]), c_int(k[1]), c_int(s[0]), c_int(s[1]), c_int(p[0]), c_int(p[1]), c_int(_dt(x)),
              c_void_p(_stream()))
        ctx.save_for_backward(idx)
        ctx.conf = (x.shape, k, s, p)
        return y

    @staticmethod
    def backward(ctx, gy):
        (idx,) = ctx.saved_tensors
        shape, k, s, p = ctx.conf
        n, c, h, w = shape
        gy = _nhwc(gy)
        ho, wo = gy.shape[2], gy.shape[3]
        dx = torch.empty(shape, dtype=gy.dtype, device=gy.device, memory_format=torch.channels_last)
        _call("mpb_maxpool_bwd", _p(gy), _p(idx), _p(dx), c_int(n), c_int(h), c_int(w), c_int(c), c_int(ho), c_int(wo),
              c_int(k[0]), c_int(k[1]), c_int(s[0]), c_int(s[1]), c_int(p[0]), c_int(p[1]), c_int(_dt(gy)),
              c_void_p(_stream()))
        return dx, None, None, None, None


def max_pool2d(x, k, s, p, ceil_mode=False):
    k, s, p = _pair(k), _pair(s), _pair(p)
    if x.dim() != 4 or x.shape[1] % 8 != 0 or x.dtype not in (torch.bfloat16, torch.float32):
        _note_fallback("max_pool2d")
        return F.max_pool2d(x, k, s, p, 1, ceil_mode)
    return _MaxPoolFn.apply(x, k, s, p, bool(ceil_mode))


class _AvgPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, s, p, ceil_mode, cip):
        x = _nhwc(x)
        n, c, h, w = x.shape
        ho, wo = _out_size(h, k[0], s[0], p[0], ceil_mode), _out_size(w, k[1], s[1], p[1], ceil_mode)
        y = torch.empty((n, c, ho, wo), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
        _call("mpb_avgpool", _p(x), _p(y), c_int(n), c_int(h), c_int(w), c_int(c), c_int(ho), c_int(wo), c_int(k[0]),
              c_int(k[1]), c_int(s[0]), c_int(s[1]), c_int(p[0]), c_int(p[1]), c_int(int(cip)), c_int(0), c_int(_dt(x)),
              c_void_p(_stream()))
        ctx.conf = (x.shape, k, s, p, cip)
        return y

    @staticmethod
    def backward(ctx, gy):
        shape, k, s, p, cip = ctx.conf
        n, c, h, w = shape
        gy = _nhwc(gy)
        dx = torch.empty(shape, dtype=gy.dtype, device=gy.device, memory_format=torch.channels_last)
        _call("mpb_avgpool", _p(gy), _p(dx), c_int(n), c_int(h), c_int(w), c_int(c), c_int(gy.shape[2]),
              c_int(gy.shape[3]), c_int(k[0]), c_int(k[1]), c_int(s[0]), c_int(s[1]), c_int(p[0]), c_int(p[1]),
              c_int(int(cip)), c_int(1), c_int(_dt(gy)), c_void_p(_stream()))
        return dx, None, None, None, None, None


def avg_pool2d(x, k, s, p, ceil_mode=False, count_include_pad=True):
    k, s, p = _pair(k), _pair(s), _pair(p)
    if x.dim() != 4 or x.shape[1] % 8 != 0 or x.dtype not in (torch.bfloat16, torch.float32):
        _note_fallback("avg_pool2d")
        return F.avg_pool2d(x, k, s, p, ceil_mode, count_include_pad)
    return _AvgPoolFn.apply(x, k, s, p, bool(ceil_mode), bool(count_include_pad))


class _AdaptiveAvgFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, out_hw):
        x = _nhwc(x)
        n, c, h, w = x.shape
        y = torch.empty((n, c, out_hw[0], out_hw[1]), dtype=x.dtype, device=x.device,
                        memory_format=torch.channels_last)
        _call("mpb_adaptive_avgpool", _p(x), _p(y), c_int(n), c_int(h), c_int(w), c_int(c), c_int(out_hw[0]),
              c_int(out_hw[1]), c_int(0), c_int(_dt(x)), c_void_p(_stream()))
        ctx.conf = (x.shape, out_hw)
        return y

    @staticmethod
    def backward(ctx, gy):
        shape, out_hw = ctx.conf
        n, c, h, w = shape
        gy = _nhwc(gy)
        dx = torch.empty(shape, dtype=gy.dtype, device=gy.device, memory_format=torch.channels_last)
        _call("mpb_adaptive_avgpool", _p(gy), _p(dx), c_int(n), c_int(h), c_int(w), c_int(c), c_int(out_hw[0]),
              c_int(out_hw[1]), c_int(1), c_int(_dt(gy)), c_void_p(_stream()))
        return dx, None


def adaptive_avg_pool2d(x, out):
    out = _pair(out)
    if x.dim() != 4 or x.shape[1] % 8 != 0 or x.dtype not in (torch.bfloat16, torch.float32):
        _note_fallback("adaptive_avg_pool2d")
        return F.adaptive_avg_pool2d(x, out)
    return _AdaptiveAvgFn.apply(x, out)


# =============================================================================== loss / metrics
class _CrossEntropyFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target):
        logits = logits.contiguous()
        b, c = logits.shape
        dl = torch.empty_like(logits)
        loss = torch.empty((), dtype=torch.float32, device=logits.device)
        _call("mpb_softmax_ce", _p(logits), _p(target), _p(dl), _p(loss), c_void_p(0), c_int(b), c_int(c),
              c_int(_dt(logits)), c_void_p(_stream()), launches=2)
        ctx.save_for_backward(dl)
        return loss

    @staticmethod
    def backward(ctx, g):
        (dl,) = ctx.saved_tensors
        return dl * g.to(dl.dtype), None


def cross_entropy(logits, target):
    if logits.dim() != 2 or logits.dtype not in (torch.bfloat16, torch.float32):
        return F.cross_entropy(logits.float(), target)
    return _CrossEntropyFn.apply(logits, target.contiguous())


def count_correct(logits, target):
    logits = logits.contiguous()
    if logits.dtype not in (torch.bfloat16, torch.float32):
        logits = logits.float()
    out = torch.empty((), dtype=torch.int64, device=logits.device)
    _call("mpb_count_correct", _p(logits), _p(target.contiguous()), _p(out), c_int(logits.shape[0]),
          c_int(logits.shape[1]), c_int(_dt(logits)), c_void_p(_stream()), launches=2)
    return out


# =============================================================================== optimizer (local half of K1)
def adam_flat(p, g, m, v, shadow, step, lr, b1, b2, eps, wd, grad_scale=1.0):
    _call("mpb_adam_flat", _p(p), _p(g), _p(m), _p(v), _p(shadow), c_ll(p.numel()), c_int(step), c_float(lr), c_float(b1),
          c_float(b2), c_float(eps), c_float(wd), c_float(grad_scale), c_void_p(_stream()))


def sgd_flat(p, g, buf, shadow, step, lr, momentum, wd, grad_scale=1.0):
    _call("mpb_sgd_flat", _p(p), _p(g), _p(buf), _p(shadow), c_ll(p.numel()), c_int(step), c_float(lr), c_float(momentum),
          c_float(wd), c_float(grad_scale), c_void_p(_stream()))


def cast_f32_bf16(src, dst):
    _call("mpb_cast_f32_bf16", _p(src), _p(dst), c_ll(src.numel()), c_void_p(_stream()))


# =============================================================================== preprocessing / data
def resize_normalize(img_u8, out_h, out_w, mode="bilinear", dtype=torch.float32, channels_last=False):
    single = img_u8.dim() == 3
    x = (img_u8.unsqueeze(0) if single else img_u8).contiguous()
    n, hi, wi, _ = x.shape
    if channels_last:
        out = torch.empty((n, 3, out_h, out_w), dtype=dtype, device=x.device, memory_format=torch.channels_last)
    else:
        out = torch.empty((n, 3, out_h, out_w), dtype=dtype, device=x.device)
    kmode = 2 if mode == "bilinear" else 1
    _call("mpb_resize", _p(x), _p(out), c_int(n), c_int(hi), c_int(wi), c_int(out_h), c_int(out_w),
          c_int(int(mode == "bicubic")), c_int(kmode), c_int(int(channels_last)), c_int(_dt(out)), c_void_p(_stream()))
    return out[0] if single else out


def resize_u8(img_u8_nhwc, out_h, out_w, mode="bicubic"):
    x = img_u8_nhwc.contiguous()
    n, hi, wi, _ = x.shape
    out = torch.empty((n, out_h, out_w, 3), dtype=torch.uint8, device=x.device)
    _call("mpb_resize", _p(x), _p(out), c_int(n), c_int(hi), c_int(wi), c_int(out_h), c_int(out_w),
          c_int(int(mode == "bicubic")), c_int(0), c_int(1), c_int(0), c_void_p(_stream()))
    return out


def normalize_u8(img_u8_nhwc, dtype=torch.float32):
    x = img_u8_nhwc.contiguous()
    n, h, w, _ = x.shape
    out = torch.empty((n, 3, h, w), dtype=dtype, device=x.device)
    _call("mpb_normalize_u8", _p(x), _p(out), c_int(n), c_int(h), c_int(w), c_int(0), c_int(_dt(out)), c_void_p(_stream()))
    return out


def synth_batch(seed, index, c, h, w, dtype, channels_last):
    n = index.numel()
    if channels_last:
        out = torch.empty((n, c, h, w), dtype=dtype, device=index.device, memory_format=torch.channels_last)
    else:
        out = torch.empty((n, c, h, w), dtype=dtype, device=index.device)
    _call("mpb_synth_images", _p(out), _p(index.contiguous()), c_int(n), c_int(c), c_int(h), c_int(w), c_ll(int(seed)),
          c_int(int(channels_last)), c_int(_dt(out)), c_void_p(_stream()))
    return out
