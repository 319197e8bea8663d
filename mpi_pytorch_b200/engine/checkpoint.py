"""Checkpoint save / resume.

File contract shared with the reference so checkpoints interchange in both directions
(SURVEY §3.5): ``<checkpoint_dir>/checkpoint_<model_name>.pt`` holding
``{'epoch', 'state_dict', 'optimizer', 'loss'}`` (/root/reference/main.py:163-168,
helpers.py:4-15); ``state_dict`` uses torchvision key names; every tensor is **CPU fp32**
because the reference loads with no ``map_location`` and a strict ``load_state_dict``;
``optimizer`` is a genuine ``torch.optim.Adam.state_dict()`` layout.

Improvements over the reference: the write is atomic (tmp file + rename), ``is_best``
actually copies to ``best_model_dir``, loading maps to the caller's device, and the
stored epoch is returned *and meant to be used* (SURVEY D5).
"""
from __future__ import annotations

import os
import shutil
from typing import Any, Dict, Tuple

import torch


def _to_cpu_fp32(obj: Any) -> Any:
    if torch.is_tensor(obj):
        t = obj.detach().to("cpu")
        if t.is_floating_point():
            t = t.to(torch.float32)
        return t.contiguous()
    if isinstance(obj, dict):
        return {k: _to_cpu_fp32(v) for k, v in obj.items()}
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_cpu_fp32(v) for v in obj)
    return obj


def make_checkpoint(epoch: int, model: torch.nn.Module, optimizer: Any, loss: float) -> Dict[str, Any]:
    """The dict of main.py:163-168, normalised to CPU fp32."""
    return {"epoch": int(epoch),
            "state_dict": _to_cpu_fp32(dict(model.state_dict())),
            "optimizer": _to_cpu_fp32(optimizer.state_dict()),
            "loss": float(loss)}


def checkpoint_path(checkpoint_dir: str, model_name: str) -> str:
    return os.path.join(checkpoint_dir, "checkpoint_{}.pt".format(model_name))


def save_checkpoint(state: Dict[str, Any], epoch: int, model_name: str, checkpoint_dir: str,
                    best_model_dir: str | None = None, is_best: bool = True) -> str:
    """Same positional signature as helpers.py:4.  Returns the path written."""
    os.makedirs(checkpoint_dir, exist_ok=True)
    path = checkpoint_path(checkpoint_dir, model_name)
    tmp = path + ".tmp.{}".format(os.getpid())
    torch.save(_to_cpu_fp32(state), tmp)
    os.replace(tmp, path)
    if best_model_dir and is_best:
        try:
            os.makedirs(best_model_dir, exist_ok=True)
            shutil.copyfile(path, os.path.join(best_model_dir, "best_{}.pt".format(model_name)))
        except OSError:
            pass
    return path


def load_checkpoint(checkpoint_fpath: str, model: torch.nn.Module, optimizer: Any = None,
                    map_location: Any = "cpu") -> Tuple[torch.nn.Module, Any, int]:
    """Same signature / return value as helpers.py:10-15."""
    # tensors, numbers, strings, dicts and lists only -- which is all the checkpoint contract contains -- so the file is
    # read with the restricted unpickler; MPB_TRUST_CHECKPOINT=1 allows arbitrary pickles (files you wrote yourself)
    trust = os.environ.get("MPB_TRUST_CHECKPOINT", "0") == "1"
    ck = torch.load(checkpoint_fpath, map_location=map_location, weights_only=not trust)
    model.load_state_dict(ck["state_dict"])
    if optimizer is not None and ck.get("optimizer") is not None:
        optimizer.load_state_dict(ck["optimizer"])
    return model, optimizer, int(ck.get("epoch", 0))
