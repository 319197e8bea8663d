"""ResNet-18 / ResNet-34 (BasicBlock family).

Module tree and registration order reproduce torchvision's ``resnet18/34`` (what
/root/reference/models.py:30-45 instantiates) so ``state_dict`` keys, shapes and the
default initialisation sequence are identical.  The forward pass is written around the
fused unit ``conv -> BN -> (+identity) -> ReLU`` (:func:`layers.conv_bn_act`) which on
B200 is one implicit-GEMM launch whose epilogue feeds the BN statistics, followed by
one fused normalise/add/ReLU pass.
"""
from __future__ import annotations

from typing import List, Optional

import torch
import torch.nn as nn

from .. import ops
from .layers import AdaptiveAvgPool2d, BatchNorm2d, Conv2d, Linear, MaxPool2d, ReLU, conv_bn_act


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes: int, planes: int, stride: int = 1, downsample: Optional[nn.Module] = None):
        super().__init__()
        self.conv1 = Conv2d(inplanes, planes, 3, stride, 1, bias=False)
        self.bn1 = BatchNorm2d(planes)
        self.relu = ReLU(inplace=True)
        self.conv2 = Conv2d(planes, planes, 3, 1, 1, bias=False)
        self.bn2 = BatchNorm2d(planes)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        identity = x
        out = conv_bn_act(x, self.conv1, self.bn1, relu=True)
        if self.downsample is not None:
            identity = conv_bn_act(x, self.downsample[0], self.downsample[1], relu=False)
        return conv_bn_act(out, self.conv2, self.bn2, relu=True, residual=identity)


class ResNet(nn.Module):
    def __init__(self, layers: List[int], num_classes: int = 1000):
        super().__init__()
        self.inplanes = 64
        self.conv1 = Conv2d(3, 64, 7, 2, 3, bias=False)
        self.bn1 = BatchNorm2d(64)
        self.relu = ReLU(inplace=True)
        self.maxpool = MaxPool2d(3, 2, 1)
        self.layer1 = self._make_layer(64, layers[0], 1)
        self.layer2 = self._make_layer(128, layers[1], 2)
        self.layer3 = self._make_layer(256, layers[2], 2)
        self.layer4 = self._make_layer(512, layers[3], 2)
        self.avgpool = AdaptiveAvgPool2d((1, 1))
        self.fc = Linear(512, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)

    def _make_layer(self, planes: int, blocks: int, stride: int) -> nn.Sequential:
        downsample = None
        if stride != 1 or self.inplanes != planes:
            downsample = nn.Sequential(Conv2d(self.inplanes, planes, 1, stride, bias=False), BatchNorm2d(planes))
        seq = [BasicBlock(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes
        for _ in range(1, blocks):
            seq.append(BasicBlock(planes, planes))
        return nn.Sequential(*seq)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = conv_bn_act(x, self.conv1, self.bn1, relu=True)
        x = self.maxpool(x)
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        x = self.avgpool(x)
        x = torch.flatten(x, 1)
        return self.fc(x)


def resnet18(num_classes: int = 1000) -> ResNet:
    return ResNet([2, 2, 2, 2], num_classes)


def resnet34(num_classes: int = 1000) -> ResNet:
    return ResNet([3, 4, 6, 3], num_classes)
