"""Inception-v3 with the auxiliary classifier.

Same attribute names as torchvision's ``inception_v3`` (``Mixed_6c.branch7x7dbl_4.bn`` ...)
so the 580-entry ``state_dict`` matches what /root/reference/models.py:83-95 builds.
Two reference defects are handled here instead of reproduced: the training
forward returns ``(logits, aux_logits)`` that the trainer combines as
``CE(logits) + 0.4*CE(aux)``, and the aux head is skipped (``None``) when the input is
too small for its 5x5 convolution (e.g. the 128x128 default image size).
"""
from __future__ import annotations

from typing import NamedTuple, Optional

import torch
import torch.nn as nn

from .. import ops
from .layers import AdaptiveAvgPool2d, BatchNorm2d, Conv2d, Dropout, Linear, MaxPool2d


class InceptionOutputs(NamedTuple):
    logits: torch.Tensor
    aux_logits: Optional[torch.Tensor]


class BasicConv2d(nn.Module):
    def __init__(self, c_in: int, c_out: int, **kw):
        super().__init__()
        self.conv = Conv2d(c_in, c_out, bias=False, **kw)
        self.bn = BatchNorm2d(c_out, eps=0.001)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return ops.conv_bn_act(x, self.conv, self.bn, relu=True)


def _avg3(x):
    return ops.avg_pool2d(x, 3, 1, 1)


class InceptionA(nn.Module):
    def __init__(self, c_in: int, pool_features: int):
        super().__init__()
        self.branch1x1 = BasicConv2d(c_in, 64, kernel_size=1)
        self.branch5x5_1 = BasicConv2d(c_in, 48, kernel_size=1)
        self.branch5x5_2 = BasicConv2d(48, 64, kernel_size=5, padding=2)
        self.branch3x3dbl_1 = BasicConv2d(c_in, 64, kernel_size=1)
        self.branch3x3dbl_2 = BasicConv2d(64, 96, kernel_size=3, padding=1)
        self.branch3x3dbl_3 = BasicConv2d(96, 96, kernel_size=3, padding=1)
        self.branch_pool = BasicConv2d(c_in, pool_features, kernel_size=1)

    def forward(self, x):
        return ops.cat_channels([
            self.branch1x1(x),
            self.branch5x5_2(self.branch5x5_1(x)),
            self.branch3x3dbl_3(self.branch3x3dbl_2(self.branch3x3dbl_1(x))),
            self.branch_pool(_avg3(x)),
        ])


class InceptionB(nn.Module):
    def __init__(self, c_in: int):
        super().__init__()
        self.branch3x3 = BasicConv2d(c_in, 384, kernel_size=3, stride=2)
        self.branch3x3dbl_1 = BasicConv2d(c_in, 64, kernel_size=1)
        self.branch3x3dbl_2 = BasicConv2d(64, 96, kernel_size=3, padding=1)
        self.branch3x3dbl_3 = BasicConv2d(96, 96, kernel_size=3, stride=2)

    def forward(self, x):
        return ops.cat_channels([
            self.branch3x3(x),
            self.branch3x3dbl_3(self.branch3x3dbl_2(self.branch3x3dbl_1(x))),
            ops.max_pool2d(x, 3, 2),
        ])


class InceptionC(nn.Module):
    def __init__(self, c_in: int, c7: int):
        super().__init__()
        self.branch1x1 = BasicConv2d(c_in, 192, kernel_size=1)
        self.branch7x7_1 = BasicConv2d(c_in, c7, kernel_size=1)
        self.branch7x7_2 = BasicConv2d(c7, c7, kernel_size=(1, 7), padding=(0, 3))
        self.branch7x7_3 = BasicConv2d(c7, 192, kernel_size=(7, 1), padding=(3, 0))
        self.branch7x7dbl_1 = BasicConv2d(c_in, c7, kernel_size=1)
        self.branch7x7dbl_2 = BasicConv2d(c7, c7, kernel_size=(7, 1), padding=(3, 0))
        self.branch7x7dbl_3 = BasicConv2d(c7, c7, kernel_size=(1, 7), padding=(0, 3))
        self.branch7x7dbl_4 = BasicConv2d(c7, c7, kernel_size=(7, 1), padding=(3, 0))
        self.branch7x7dbl_5 = BasicConv2d(c7, 192, kernel_size=(1, 7), padding=(0, 3))
        self.branch_pool = BasicConv2d(c_in, 192, kernel_size=1)

    def forward(self, x):
        d = self.branch7x7dbl_1(x)
        d = self.branch7x7dbl_5(self.branch7x7dbl_4(self.branch7x7dbl_3(self.branch7x7dbl_2(d))))
        return ops.cat_channels([
            self.branch1x1(x),
            self.branch7x7_3(self.branch7x7_2(self.branch7x7_1(x))),
            d,
            self.branch_pool(_avg3(x)),
        ])


class InceptionD(nn.Module):
    def __init__(self, c_in: int):
        super().__init__()
        self.branch3x3_1 = BasicConv2d(c_in, 192, kernel_size=1)
        self.branch3x3_2 = BasicConv2d(192, 320, kernel_size=3, stride=2)
        self.branch7x7x3_1 = BasicConv2d(c_in, 192, kernel_size=1)
        self.branch7x7x3_2 = BasicConv2d(192, 192, kernel_size=(1, 7), padding=(0, 3))
        self.branch7x7x3_3 = BasicConv2d(192, 192, kernel_size=(7, 1), padding=(3, 0))
        self.branch7x7x3_4 = BasicConv2d(192, 192, kernel_size=3, stride=2)

    def forward(self, x):
        return ops.cat_channels([
            self.branch3x3_2(self.branch3x3_1(x)),
            self.branch7x7x3_4(self.branch7x7x3_3(self.branch7x7x3_2(self.branch7x7x3_1(x)))),
            ops.max_pool2d(x, 3, 2),
        ])


class InceptionE(nn.Module):
    def __init__(self, c_in: int):
        super().__init__()
        self.branch1x1 = BasicConv2d(c_in, 320, kernel_size=1)
        self.branch3x3_1 = BasicConv2d(c_in, 384, kernel_size=1)
        self.branch3x3_2a = BasicConv2d(384, 384, kernel_size=(1, 3), padding=(0, 1))
        self.branch3x3_2b = BasicConv2d(384, 384, kernel_size=(3, 1), padding=(1, 0))
        self.branch3x3dbl_1 = BasicConv2d(c_in, 448, kernel_size=1)
        self.branch3x3dbl_2 = BasicConv2d(448, 384, kernel_size=3, padding=1)
        self.branch3x3dbl_3a = BasicConv2d(384, 384, kernel_size=(1, 3), padding=(0, 1))
        self.branch3x3dbl_3b = BasicConv2d(384, 384, kernel_size=(3, 1), padding=(1, 0))
        self.branch_pool = BasicConv2d(c_in, 192, kernel_size=1)

    def forward(self, x):
        b3 = self.branch3x3_1(x)
        bd = self.branch3x3dbl_2(self.branch3x3dbl_1(x))
        return ops.cat_channels([
            self.branch1x1(x),
            ops.cat_channels([self.branch3x3_2a(b3), self.branch3x3_2b(b3)]),
            ops.cat_channels([self.branch3x3dbl_3a(bd), self.branch3x3dbl_3b(bd)]),
            self.branch_pool(_avg3(x)),
        ])


class InceptionAux(nn.Module):
    def __init__(self, c_in: int, num_classes: int):
        super().__init__()
        self.conv0 = BasicConv2d(c_in, 128, kernel_size=1)
        self.conv1 = BasicConv2d(128, 768, kernel_size=5)
        self.conv1.stddev = 0.01  # type: ignore[assignment]
        self.fc = Linear(768, num_classes)
        self.fc.stddev = 0.001  # type: ignore[assignment]

    def forward(self, x: torch.Tensor) -> Optional[torch.Tensor]:
        if x.shape[-1] < 5 or x.shape[-2] < 5:
            return None
        x = ops.avg_pool2d(x, 5, 3)
        if x.shape[-1] < 5 or x.shape[-2] < 5:
            return None
        x = self.conv1(self.conv0(x))
        x = ops.adaptive_avg_pool2d(x, (1, 1))
        return self.fc(torch.flatten(x, 1))


class Inception3(nn.Module):
    def __init__(self, num_classes: int = 1000, aux_logits: bool = True, dropout: float = 0.5):
        super().__init__()
        self.aux_logits = aux_logits
        self.transform_input = False
        self.Conv2d_1a_3x3 = BasicConv2d(3, 32, kernel_size=3, stride=2)
        self.Conv2d_2a_3x3 = BasicConv2d(32, 32, kernel_size=3)
        self.Conv2d_2b_3x3 = BasicConv2d(32, 64, kernel_size=3, padding=1)
        self.maxpool1 = MaxPool2d(3, 2)
        self.Conv2d_3b_1x1 = BasicConv2d(64, 80, kernel_size=1)
        self.Conv2d_4a_3x3 = BasicConv2d(80, 192, kernel_size=3)
        self.maxpool2 = MaxPool2d(3, 2)
        self.Mixed_5b = InceptionA(192, 32)
        self.Mixed_5c = InceptionA(256, 64)
        self.Mixed_5d = InceptionA(288, 64)
        self.Mixed_6a = InceptionB(288)
        self.Mixed_6b = InceptionC(768, 128)
        self.Mixed_6c = InceptionC(768, 160)
        self.Mixed_6d = InceptionC(768, 160)
        self.Mixed_6e = InceptionC(768, 192)
        self.AuxLogits: Optional[nn.Module] = InceptionAux(768, num_classes) if aux_logits else None
        self.Mixed_7a = InceptionD(768)
        self.Mixed_7b = InceptionE(1280)
        self.Mixed_7c = InceptionE(2048)
        self.avgpool = AdaptiveAvgPool2d((1, 1))
        self.dropout = Dropout(p=dropout)
        self.fc = Linear(2048, num_classes)
        for m in self.modules():
            if isinstance(m, (nn.Conv2d, nn.Linear)):
                std = float(getattr(m, "stddev", 0.1))
                nn.init.trunc_normal_(m.weight, mean=0.0, std=std, a=-2, b=2)
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)

    def forward(self, x: torch.Tensor):
        x = self.Conv2d_2b_3x3(self.Conv2d_2a_3x3(self.Conv2d_1a_3x3(x)))
        x = self.maxpool1(x)
        x = self.Conv2d_4a_3x3(self.Conv2d_3b_1x1(x))
        x = self.maxpool2(x)
        x = self.Mixed_5d(self.Mixed_5c(self.Mixed_5b(x)))
        x = self.Mixed_6a(x)
        x = self.Mixed_6e(self.Mixed_6d(self.Mixed_6c(self.Mixed_6b(x))))
        aux = None
        if self.AuxLogits is not None and self.training:
            aux = self.AuxLogits(x)
        x = self.Mixed_7c(self.Mixed_7b(self.Mixed_7a(x)))
        x = self.dropout(self.avgpool(x))
        logits = self.fc(torch.flatten(x, 1))
        if self.training and self.aux_logits:
            return InceptionOutputs(logits, aux)
        return logits


def inception_v3(num_classes: int = 1000) -> Inception3:
    return Inception3(num_classes, aux_logits=True)
