"""Model zoo factory.

Same contract as the reference's ``initialize_model(model_name, num_classes,
feature_extract, use_pretrained=True) -> (model, input_size)``
(/root/reference/models.py:16-101): seven names, the classifier head replaced by a
fresh layer with ``num_classes`` outputs, the returned ``input_size`` per architecture
(224; 128 for resnet34 as the reference has it; 299 for inception).  ``vgg`` means
VGG-11-bn exactly as in the reference; ``vgg16`` is an *additional* name for
BASELINE.json's all-reduce-bound configuration.

Differences, all deliberate (SURVEY §4 defects):
* an unknown name raises ``ValueError`` instead of ``print`` + ``exit()`` (models.py:97-99);
* ``use_pretrained`` is honoured but weights are only taken from the local torch hub
  cache -- there is no network on the build/GPU boxes, so the default falls back to
  random initialisation with a warning instead of crashing.
"""
from __future__ import annotations

import os
import warnings
from typing import Callable, Dict, Tuple

import torch
import torch.nn as nn

from . import densenet as _densenet
from . import inception as _inception
from . import plain as _plain
from . import resnet as _resnet
from .layers import Conv2d, Linear

# name -> (constructor, input_size returned by the reference, torchvision weight file stem)
_ZOO: Dict[str, Tuple[Callable[..., nn.Module], int, str]] = {
    "resnet18": (_resnet.resnet18, 224, "resnet18"),
    "resnet34": (_resnet.resnet34, 128, "resnet34"),     # models.py:45 returns 128 for resnet34
    "alexnet": (_plain.alexnet, 224, "alexnet"),
    "vgg": (_plain.vgg11_bn, 224, "vgg11_bn"),
    "vgg16": (_plain.vgg16, 224, "vgg16"),
    "squeezenet": (_plain.squeezenet1_0, 224, "squeezenet1_0"),
    "densenet": (_densenet.densenet121, 224, "densenet121"),
    "inception": (_inception.inception_v3, 299, "inception_v3"),
}

MODEL_NAMES = tuple(_ZOO)


def set_parameter_requires_grad(model: nn.Module, feature_extracting: bool) -> None:
    """Freeze every existing parameter when feature extracting (models.py:5-13)."""
    if feature_extracting:
        for p in model.parameters():
            p.requires_grad_(False)


def _try_load_pretrained(model: nn.Module, stem: str) -> bool:
    hub = os.path.join(torch.hub.get_dir(), "checkpoints")
    if os.path.isdir(hub):
        for fn in sorted(os.listdir(hub)):
            if fn.startswith(stem + "-") and fn.endswith(".pth"):
                sd = torch.load(os.path.join(hub, fn), map_location="cpu")
                model.load_state_dict(sd, strict=False)
                return True
    return False


def _replace_head(name: str, model: nn.Module, num_classes: int) -> None:
    if name in ("resnet18", "resnet34"):
        model.fc = Linear(model.fc.in_features, num_classes)
    elif name in ("alexnet", "vgg", "vgg16"):
        model.classifier[6] = Linear(model.classifier[6].in_features, num_classes)
    elif name == "squeezenet":
        model.classifier[1] = Conv2d(512, num_classes, kernel_size=(1, 1), stride=(1, 1))
        model.num_classes = num_classes
    elif name == "densenet":
        model.classifier = Linear(model.classifier.in_features, num_classes)
    elif name == "inception":
        model.AuxLogits.fc = Linear(model.AuxLogits.fc.in_features, num_classes)
        model.fc = Linear(model.fc.in_features, num_classes)


def initialize_model(model_name: str, num_classes: int, feature_extract: bool,
                     use_pretrained: bool = True) -> Tuple[nn.Module, int]:
    if model_name not in _ZOO:
        raise ValueError("Invalid model name {!r}; choose one of {}".format(model_name, ", ".join(MODEL_NAMES)))
    ctor, input_size, stem = _ZOO[model_name]
    model = ctor()                                   # ImageNet-shaped (1000-way) body, like torchvision
    if use_pretrained and not _try_load_pretrained(model, stem):
        warnings.warn("use_pretrained=True but no '{}-*.pth' in the local torch hub cache and there is no "
                      "network: continuing with random initialisation".format(stem), stacklevel=2)
    set_parameter_requires_grad(model, feature_extract)
    _replace_head(model_name, model, num_classes)    # created after the freeze -> stays trainable
    return model, input_size
