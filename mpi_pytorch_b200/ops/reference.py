"""Plain-torch implementations of every op in :mod:`mpi_pytorch_b200.ops`.

These are (a) the CPU execution path (tests, the gloo multi-process paths) and (b) the
numerical oracle each CUDA kernel is tested against.  They mirror what the reference
reaches through torchvision/ATen.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn.functional as F


def conv2d(x, w, b, stride, padding, dilation=(1, 1), groups=1):
    return F.conv2d(x, w.to(x.dtype), None if b is None else b.to(x.dtype), stride, padding, dilation, groups)


def linear(x, w, b):
    return F.linear(x, w.to(x.dtype), None if b is None else b.to(x.dtype))


def batch_norm_act(x, bn, residual=None, relu=False):
    if bn.training and bn.track_running_stats and bn.num_batches_tracked is not None:
        bn.num_batches_tracked.add_(1)
    use_batch = bn.training or not bn.track_running_stats
    momentum = 0.0 if bn.momentum is None else bn.momentum
    if bn.momentum is None and bn.training and bn.track_running_stats:
        momentum = 1.0 / float(bn.num_batches_tracked)
    y = F.batch_norm(x, bn.running_mean if (not bn.training or bn.track_running_stats) else None,
                     bn.running_var if (not bn.training or bn.track_running_stats) else None,
                     bn.weight, bn.bias, use_batch, momentum, bn.eps)
    if residual is not None:
        y = y + residual
    if relu:
        y = F.relu(y)
    return y


def relu(x, inplace=False):
    return F.relu(x, inplace=inplace)


def max_pool2d(x, k, s, p, ceil_mode=False):
    return F.max_pool2d(x, k, s, p, 1, ceil_mode)


def avg_pool2d(x, k, s, p, ceil_mode=False, count_include_pad=True):
    return F.avg_pool2d(x, k, s, p, ceil_mode, count_include_pad)


def adaptive_avg_pool2d(x, out):
    return F.adaptive_avg_pool2d(x, out)


def dropout(x, p, training, inplace=False):
    return F.dropout(x, p, training, inplace)


def cross_entropy(logits, target):
    return F.cross_entropy(logits.float(), target)


def count_correct(logits, target) -> torch.Tensor:
    """argmax + compare + sum (main.py:182-183, evaluation_pipeline.py:157-158) as one int64."""
    return (logits.argmax(dim=1) == target).sum().to(torch.int64)


_MEAN = (0.485, 0.456, 0.406)
_STD = (0.229, 0.224, 0.225)


def resize_normalize(img_u8_hwc: torch.Tensor, out_h: int, out_w: int, mode: str = "bilinear",
                     dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """uint8 HWC (or NHWC) -> resized -> (x/255-mean)/std -> CHW (NCHW).

    ``bilinear`` = train-time semantics (ToTensor -> Resize on the tensor -> Normalize,
    /root/reference/main.py:62-65; torchvision's tensor Resize antialiases by default);
    ``bicubic`` = eval-pipeline semantics (PIL resize before ToTensor,
    evaluation_pipeline.py:89,116-122)."""
    single = img_u8_hwc.dim() == 3
    x = img_u8_hwc.unsqueeze(0) if single else img_u8_hwc
    x = x.permute(0, 3, 1, 2).float()
    if mode == "bilinear":
        x = F.interpolate(x / 255.0, size=(out_h, out_w), mode="bilinear", align_corners=False, antialias=True)
    elif mode == "bicubic":
        # PIL resizes uint8 and rounds back to uint8 before ToTensor
        x = F.interpolate(x, size=(out_h, out_w), mode="bicubic", align_corners=False, antialias=True)
        x = x.round().clamp(0, 255) / 255.0
    else:
        raise ValueError(mode)
    mean = torch.tensor(_MEAN, device=x.device).view(1, 3, 1, 1)
    std = torch.tensor(_STD, device=x.device).view(1, 3, 1, 1)
    x = ((x - mean) / std).to(dtype)
    return x[0] if single else x


def resize_u8(img_u8_nhwc: torch.Tensor, out_h: int, out_w: int, mode: str = "bicubic") -> torch.Tensor:
    """uint8 NHWC -> uint8 NHWC resize (pipeline stage 1, evaluation_pipeline.py:89; PIL's default
    filter for the pinned Pillow is bicubic with antialiasing, result rounded back to uint8)."""
    x = img_u8_nhwc.permute(0, 3, 1, 2).float()
    x = F.interpolate(x, size=(out_h, out_w), mode=mode, align_corners=False, antialias=True)
    return x.round().clamp(0, 255).to(torch.uint8).permute(0, 2, 3, 1).contiguous()


def normalize_u8(img_u8_nhwc: torch.Tensor, dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """uint8 NHWC -> ToTensor + Normalize(ImageNet) -> NCHW (pipeline stage 2, evaluation_pipeline.py:116-122)."""
    x = img_u8_nhwc.permute(0, 3, 1, 2).float() / 255.0
    mean = torch.tensor(_MEAN, device=x.device).view(1, 3, 1, 1)
    std = torch.tensor(_STD, device=x.device).view(1, 3, 1, 1)
    return ((x - mean) / std).to(dtype).contiguous()


def adam_step(p, g, m, v, step: int, lr, b1=0.9, b2=0.999, eps=1e-8, wd=0.0, grad_scale=1.0):
    """torch.optim.Adam (non-amsgrad) single-tensor math on flat tensors, in place."""
    g = g.float() * grad_scale
    if wd:
        g = g + wd * p
    m.mul_(b1).add_(g, alpha=1 - b1)
    v.mul_(b2).addcmul_(g, g, value=1 - b2)
    bc1 = 1 - b1 ** step
    bc2 = 1 - b2 ** step
    denom = (v.sqrt() / (bc2 ** 0.5)).add_(eps)
    p.addcdiv_(m, denom, value=-lr / bc1)


def sgd_step(p, g, buf, step: int, lr, momentum=0.9, wd=0.0, grad_scale=1.0):
    g = g.float() * grad_scale
    if wd:
        g = g + wd * p
    if momentum:
        if step == 1:
            buf.copy_(g)
        else:
            buf.mul_(momentum).add_(g)
        g = buf
    p.add_(g, alpha=-lr)
