"""Tensor-core (tcgen05 / TMEM / TMA) GEMM and implicit-GEMM convolution front end.

``*_supported`` decide whether a call can run on the hand-written sm_100a kernels in
``csrc/tc_gemm.cu``; anything else is counted as a fallback by ``ops.native`` and runs on the
library kernel.  Filled in by the K3-K6 work (see DESIGN.md for the status table).
"""
from __future__ import annotations

import os

import torch

ENABLED = os.environ.get("MPB_TC", "1").lower() not in ("0", "false", "off")


def gemm_supported(x: torch.Tensor, w: torch.Tensor) -> bool:
    return False


def conv_supported(x, w, stride, padding, dilation, groups) -> bool:
    return False
