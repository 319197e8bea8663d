"""DenseNet-121.

Names (``features.denseblock2.denselayer7.norm1`` ...) and construction order match
torchvision's ``densenet121`` so the 727-entry ``state_dict`` interchanges with what
/root/reference/models.py:74-81 builds.  BN -> ReLU pairs run as one fused
normalise+activate kernel; concatenation goes through ``ops.cat_channels``.
"""
from __future__ import annotations

from collections import OrderedDict
from typing import List, Tuple

import torch
import torch.nn as nn

from .. import ops
from .layers import AvgPool2d, BatchNorm2d, Conv2d, Linear, MaxPool2d, ReLU


class _DenseLayer(nn.Module):
    def __init__(self, c_in: int, growth: int, bn_size: int):
        super().__init__()
        self.norm1 = BatchNorm2d(c_in)
        self.relu1 = ReLU(inplace=True)
        self.conv1 = Conv2d(c_in, bn_size * growth, 1, 1, bias=False)
        self.norm2 = BatchNorm2d(bn_size * growth)
        self.relu2 = ReLU(inplace=True)
        self.conv2 = Conv2d(bn_size * growth, growth, 3, 1, 1, bias=False)

    def forward(self, feats: List[torch.Tensor]) -> torch.Tensor:
        x = ops.cat_channels(feats) if len(feats) > 1 else feats[0]
        x = self.conv1(self.norm1(x, relu=True))
        return self.conv2(self.norm2(x, relu=True))


class _DenseBlock(nn.ModuleDict):
    def __init__(self, n_layers: int, c_in: int, bn_size: int, growth: int):
        super().__init__()
        for i in range(n_layers):
            self.add_module("denselayer%d" % (i + 1), _DenseLayer(c_in + i * growth, growth, bn_size))

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        feats = [x]
        for _, layer in self.items():
            feats.append(layer(feats))
        return ops.cat_channels(feats)


class _Transition(nn.Sequential):
    def __init__(self, c_in: int, c_out: int):
        super().__init__()
        self.norm = BatchNorm2d(c_in)
        self.relu = ReLU(inplace=True)
        self.conv = Conv2d(c_in, c_out, 1, 1, bias=False)
        self.pool = AvgPool2d(2, 2)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        return self.pool(self.conv(self.norm(x, relu=True)))


class DenseNet(nn.Module):
    def __init__(self, growth: int = 32, blocks: Tuple[int, ...] = (6, 12, 24, 16), init_features: int = 64,
                 bn_size: int = 4, num_classes: int = 1000):
        super().__init__()
        self.features = nn.Sequential(OrderedDict([
            ("conv0", Conv2d(3, init_features, 7, 2, 3, bias=False)),
            ("norm0", BatchNorm2d(init_features)),
            ("relu0", ReLU(inplace=True)),
            ("pool0", MaxPool2d(3, 2, 1)),
        ]))
        c = init_features
        for i, n in enumerate(blocks):
            self.features.add_module("denseblock%d" % (i + 1), _DenseBlock(n, c, bn_size, growth))
            c += n * growth
            if i != len(blocks) - 1:
                self.features.add_module("transition%d" % (i + 1), _Transition(c, c // 2))
                c //= 2
        self.features.add_module("norm5", BatchNorm2d(c))
        self.classifier = Linear(c, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight)
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)
            elif isinstance(m, nn.Linear):
                nn.init.constant_(m.bias, 0)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        f = self.features
        x = ops.conv_bn_act(x, f.conv0, f.norm0, relu=True)
        x = f.pool0(x)
        for name, mod in f.named_children():
            if name.startswith("denseblock") or name.startswith("transition"):
                x = mod(x)
        x = f.norm5(x, relu=True)
        x = ops.adaptive_avg_pool2d(x, (1, 1))
        return self.classifier(torch.flatten(x, 1))


def densenet121(num_classes: int = 1000) -> DenseNet:
    return DenseNet(32, (6, 12, 24, 16), 64, 4, num_classes)
