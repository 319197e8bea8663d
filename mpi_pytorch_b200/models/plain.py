"""The three "plain stack" members of the zoo: AlexNet, VGG (11-bn / 16), SqueezeNet 1.0.

Index positions inside ``features`` / ``classifier`` follow torchvision exactly, because
the reference replaces heads by index (``classifier[6]`` for AlexNet/VGG,
``classifier[1]`` for SqueezeNet; /root/reference/models.py:47-72) and checkpoints are
keyed by those indices.
"""
from __future__ import annotations

from typing import List, Union

import torch
import torch.nn as nn
import torch.nn.init as init

from .. import ops
from .layers import AdaptiveAvgPool2d, BatchNorm2d, Conv2d, Dropout, Linear, MaxPool2d, ReLU


# ---------------------------------------------------------------------------- AlexNet
def run_stack(features: nn.Sequential, x: torch.Tensor) -> torch.Tensor:
    """Run a torchvision-style ``features`` stack, fusing Conv -> BN -> ReLU triples (statistics and activation in
    the conv / BN kernels) and Conv -> ReLU pairs (activation in the conv epilogue, mask + bias gradient in one
    backward pass) where the stack has them.  Module indices -- and so state_dict names -- are untouched."""
    mods = list(features)
    i = 0
    while i < len(mods):
        m = mods[i]
        if isinstance(m, Conv2d) and i + 2 < len(mods) and isinstance(mods[i + 1], BatchNorm2d) \
                and isinstance(mods[i + 2], ReLU):
            x = ops.conv_bn_act(x, m, mods[i + 1], relu=True)
            i += 3
        elif isinstance(m, Conv2d) and i + 1 < len(mods) and isinstance(mods[i + 1], ReLU):
            x = ops.conv_relu(x, m)
            i += 2
        else:
            x = m(x)
            i += 1
    return x


class AlexNet(nn.Module):
    def __init__(self, num_classes: int = 1000, dropout: float = 0.5):
        super().__init__()
        self.features = nn.Sequential(
            Conv2d(3, 64, 11, 4, 2), ReLU(inplace=True), MaxPool2d(3, 2),
            Conv2d(64, 192, 5, 1, 2), ReLU(inplace=True), MaxPool2d(3, 2),
            Conv2d(192, 384, 3, 1, 1), ReLU(inplace=True),
            Conv2d(384, 256, 3, 1, 1), ReLU(inplace=True),
            Conv2d(256, 256, 3, 1, 1), ReLU(inplace=True), MaxPool2d(3, 2),
        )
        self.avgpool = AdaptiveAvgPool2d((6, 6))
        self.classifier = nn.Sequential(
            Dropout(p=dropout), Linear(256 * 6 * 6, 4096), ReLU(inplace=True),
            Dropout(p=dropout), Linear(4096, 4096), ReLU(inplace=True),
            Linear(4096, num_classes),
        )

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.avgpool(run_stack(self.features, x))
        return self.classifier(torch.flatten(x, 1))


# ---------------------------------------------------------------------------- VGG
_VGG_CFG = {
    "A": [64, "M", 128, "M", 256, 256, "M", 512, 512, "M", 512, 512, "M"],
    "D": [64, 64, "M", 128, 128, "M", 256, 256, 256, "M", 512, 512, 512, "M", 512, 512, 512, "M"],
}


def _vgg_features(cfg: List[Union[str, int]], batch_norm: bool) -> nn.Sequential:
    seq: List[nn.Module] = []
    c_in = 3
    for v in cfg:
        if v == "M":
            seq.append(MaxPool2d(2, 2))
        else:
            seq.append(Conv2d(c_in, int(v), 3, padding=1))
            if batch_norm:
                seq.append(BatchNorm2d(int(v)))
            seq.append(ReLU(inplace=True))
            c_in = int(v)
    return nn.Sequential(*seq)


class VGG(nn.Module):
    def __init__(self, cfg: str, batch_norm: bool, num_classes: int = 1000, dropout: float = 0.5):
        super().__init__()
        self.features = _vgg_features(_VGG_CFG[cfg], batch_norm)
        self.avgpool = AdaptiveAvgPool2d((7, 7))
        self.classifier = nn.Sequential(
            Linear(512 * 7 * 7, 4096), ReLU(True), Dropout(p=dropout),
            Linear(4096, 4096), ReLU(True), Dropout(p=dropout),
            Linear(4096, num_classes),
        )
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
                if m.bias is not None:
                    init.constant_(m.bias, 0)
            elif isinstance(m, nn.BatchNorm2d):
                init.constant_(m.weight, 1)
                init.constant_(m.bias, 0)
            elif isinstance(m, nn.Linear):
                init.normal_(m.weight, 0, 0.01)
                init.constant_(m.bias, 0)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = run_stack(self.features, x)
        x = self.avgpool(x)
        return self.classifier(torch.flatten(x, 1))


def vgg11_bn(num_classes: int = 1000) -> VGG:
    return VGG("A", True, num_classes)


def vgg16(num_classes: int = 1000) -> VGG:
    return VGG("D", False, num_classes)


# ---------------------------------------------------------------------------- SqueezeNet
class Fire(nn.Module):
    def __init__(self, inplanes: int, squeeze: int, e1: int, e3: int):
        super().__init__()
        self.inplanes = inplanes
        self.squeeze = Conv2d(inplanes, squeeze, 1)
        self.squeeze_activation = ReLU(inplace=True)
        self.expand1x1 = Conv2d(squeeze, e1, 1)
        self.expand1x1_activation = ReLU(inplace=True)
        self.expand3x3 = Conv2d(squeeze, e3, 3, padding=1)
        self.expand3x3_activation = ReLU(inplace=True)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.squeeze_activation(self.squeeze(x))
        return ops.cat_channels([self.expand1x1_activation(self.expand1x1(x)),
                                 self.expand3x3_activation(self.expand3x3(x))])


class SqueezeNet(nn.Module):
    """SqueezeNet 1.0 (the variant /root/reference/models.py:68 builds)."""

    def __init__(self, num_classes: int = 1000, dropout: float = 0.5):
        super().__init__()
        self.num_classes = num_classes
        self.features = nn.Sequential(
            Conv2d(3, 96, 7, 2), ReLU(inplace=True), MaxPool2d(3, 2, ceil_mode=True),
            Fire(96, 16, 64, 64), Fire(128, 16, 64, 64), Fire(128, 32, 128, 128),
            MaxPool2d(3, 2, ceil_mode=True),
            Fire(256, 32, 128, 128), Fire(256, 48, 192, 192), Fire(384, 48, 192, 192), Fire(384, 64, 256, 256),
            MaxPool2d(3, 2, ceil_mode=True),
            Fire(512, 64, 256, 256),
        )
        final_conv = Conv2d(512, num_classes, 1)
        self.classifier = nn.Sequential(Dropout(p=dropout), final_conv, ReLU(inplace=True),
                                        AdaptiveAvgPool2d((1, 1)))
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                if m is final_conv:
                    init.normal_(m.weight, mean=0.0, std=0.01)
                else:
                    init.kaiming_uniform_(m.weight)
                if m.bias is not None:
                    init.constant_(m.bias, 0)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        x = self.classifier(self.features(x))
        return torch.flatten(x, 1)


def alexnet(num_classes: int = 1000) -> AlexNet:
    return AlexNet(num_classes)


def squeezenet1_0(num_classes: int = 1000) -> SqueezeNet:
    return SqueezeNet(num_classes)
