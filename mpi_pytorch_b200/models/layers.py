"""Layer modules of the model zoo.

Each class subclasses the torch module of the same name so parameter / buffer names,
shapes and default initialisation are exactly those torchvision produces (the
checkpoint contract of SURVEY §3.5: ``state_dict`` keys must interchange with
/root/reference/models.py + helpers.py).  ``forward`` goes through
:mod:`mpi_pytorch_b200.ops`, which runs the hand-written sm_100a kernels on CUDA and
a plain torch implementation on CPU.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from .. import ops


class Conv2d(nn.Conv2d):
    def forward(self, x: torch.Tensor) -> torch.Tensor:  # noqa: D401
        return ops.conv2d(x, self.weight, self.bias, self.stride, self.padding, self.dilation, self.groups)


class Linear(nn.Linear):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return ops.linear(x, self.weight, self.bias)


class BatchNorm2d(nn.BatchNorm2d):
    """BatchNorm with optional fused ReLU / residual add (``ops.batch_norm_act``)."""

    def forward(self, x: torch.Tensor, residual: Optional[torch.Tensor] = None, relu: bool = False) -> torch.Tensor:
        return ops.batch_norm_act(x, self, residual=residual, relu=relu)


class ReLU(nn.ReLU):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return ops.relu(x, inplace=self.inplace)


class MaxPool2d(nn.MaxPool2d):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return ops.max_pool2d(x, self.kernel_size, self.stride, self.padding, self.ceil_mode)


class AvgPool2d(nn.AvgPool2d):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return ops.avg_pool2d(x, self.kernel_size, self.stride, self.padding, self.ceil_mode,
                              self.count_include_pad)


class AdaptiveAvgPool2d(nn.AdaptiveAvgPool2d):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return ops.adaptive_avg_pool2d(x, self.output_size)


class Dropout(nn.Dropout):
    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return ops.dropout(x, self.p, self.training, self.inplace)


def conv_bn_act(x: torch.Tensor, conv: Conv2d, bn: BatchNorm2d, relu: bool = True,
                residual: Optional[torch.Tensor] = None) -> torch.Tensor:
    """conv -> BN(train stats or eval affine) -> (+residual) -> ReLU as one fused unit.
    On CUDA the conv epilogue accumulates the BN batch statistics (no extra pass over y)."""
    return ops.conv_bn_act(x, conv, bn, relu=relu, residual=residual)
