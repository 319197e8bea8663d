"""Functional op API of the framework.

On CUDA tensors every function runs a hand-written sm_100a kernel from
``mpi_pytorch_b200/csrc`` (through :mod:`.cuda_impl`, reached as ``ops.native()``); on CPU tensors it runs the plain
torch implementation in :mod:`.reference`, which is also the oracle the kernels are
tested against.  There is no silent CUDA fallback: if the extension is missing on a
GPU box the import of :mod:`.cuda_impl` raises (set ``MPB_NATIVE_OPS=0`` to run the
torch/cuDNN baseline path deliberately, e.g. for the DDP comparison).
"""
from __future__ import annotations

import os
from typing import Optional

import torch

from . import reference as ref

_STATE = {"enabled": os.environ.get("MPB_NATIVE_OPS", "1").lower() not in ("0", "false", "off"),
          "native": None}


def set_native(enabled: bool) -> None:
    _STATE["enabled"] = bool(enabled)


def native_enabled() -> bool:
    return bool(_STATE["enabled"])


def native():
    """The loaded extension wrapper (imports / builds it on first use)."""
    if _STATE["native"] is None:
        import importlib
        _STATE["native"] = importlib.import_module(".cuda_impl", __name__)
    return _STATE["native"]


def _nat(x: torch.Tensor) -> bool:
    return _STATE["enabled"] and x.is_cuda


# --------------------------------------------------------------------------- ops
def conv2d(x, w, b, stride, padding, dilation=(1, 1), groups=1):
    if _nat(x):
        return native().conv2d(x, w, b, stride, padding, dilation, groups)
    return ref.conv2d(x, w, b, stride, padding, dilation, groups)


def linear(x, w, b):
    if _nat(x):
        return native().linear(x, w, b)
    return ref.linear(x, w, b)


def batch_norm_act(x, bn, residual=None, relu=False):
    if _nat(x):
        return native().batch_norm_act(x, bn, residual, relu)
    return ref.batch_norm_act(x, bn, residual, relu)


def conv_bn_act(x, conv, bn, relu=True, residual=None):
    if _nat(x):
        return native().conv_bn_act(x, conv, bn, relu, residual)
    y = ref.conv2d(x, conv.weight, conv.bias, conv.stride, conv.padding, conv.dilation, conv.groups)
    return ref.batch_norm_act(y, bn, residual, relu)


def conv_relu(x, conv):
    """Conv2d followed by ReLU (VGG / AlexNet / SqueezeNet stacks, /root/reference/models.py:47-72)."""
    if _nat(x):
        return native().conv_relu(x, conv)
    return ref.relu(ref.conv2d(x, conv.weight, conv.bias, conv.stride, conv.padding, conv.dilation, conv.groups))


def relu(x, inplace=False):
    if _nat(x):
        return native().relu(x, inplace)
    return ref.relu(x, inplace)


def max_pool2d(x, k, s=None, p=0, ceil_mode=False):
    s = k if s is None else s
    if _nat(x):
        return native().max_pool2d(x, k, s, p, ceil_mode)
    return ref.max_pool2d(x, k, s, p, ceil_mode)


def avg_pool2d(x, k, s=None, p=0, ceil_mode=False, count_include_pad=True):
    s = k if s is None else s
    if _nat(x):
        return native().avg_pool2d(x, k, s, p, ceil_mode, count_include_pad)
    return ref.avg_pool2d(x, k, s, p, ceil_mode, count_include_pad)


def adaptive_avg_pool2d(x, out):
    if _nat(x):
        return native().adaptive_avg_pool2d(x, out)
    return ref.adaptive_avg_pool2d(x, out)


def dropout(x, p, training, inplace=False):
    if not training or p == 0.0:
        return x
    if _nat(x):
        return native().dropout(x, p)
    return ref.dropout(x, p, training, inplace)


def cross_entropy(logits, target):
    """Mean softmax cross-entropy (nn.CrossEntropyLoss(), /root/reference/main.py:134,150)."""
    if _nat(logits):
        return native().cross_entropy(logits, target)
    return ref.cross_entropy(logits, target)


def count_correct(logits, target):
    if _nat(logits):
        return native().count_correct(logits, target)
    return ref.count_correct(logits, target)


def resize_normalize(img_u8, out_h, out_w, mode="bilinear", dtype=torch.float32, channels_last=False):
    if _nat(img_u8):
        return native().resize_normalize(img_u8, out_h, out_w, mode, dtype, channels_last)
    out = ref.resize_normalize(img_u8, out_h, out_w, mode, dtype)
    if channels_last and out.dim() == 4:
        out = out.contiguous(memory_format=torch.channels_last)
    return out


def resize_u8(img_u8_nhwc, out_h, out_w, mode="bicubic"):
    if _nat(img_u8_nhwc):
        return native().resize_u8(img_u8_nhwc, out_h, out_w, mode)
    return ref.resize_u8(img_u8_nhwc, out_h, out_w, mode)


def normalize_u8(img_u8_nhwc, dtype=torch.float32):
    if _nat(img_u8_nhwc):
        return native().normalize_u8(img_u8_nhwc, dtype)
    return ref.normalize_u8(img_u8_nhwc, dtype)


def cat_channels(tensors):
    """Channel concat (DenseNet / Inception / Fire)."""
    if _nat(tensors[0]):
        return native().cat_channels(tensors)
    return torch.cat(tensors, dim=1)


def prepare_inference(model, dtype=torch.bfloat16):
    """Forward-only preparation of a model outside the DP engine (bf16 compute copies of the GEMM-shaped weights on
    CUDA; a no-op besides ``eval()`` on the torch path)."""
    if _STATE["enabled"] and any(p.is_cuda for p in model.parameters()):
        return native().prepare_inference(model, dtype)
    return model.eval()
