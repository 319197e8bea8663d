"""Run configuration.

The reference configures a run by editing module-level constants in ``utils.py``
(/root/reference/utils.py:4-45) and reading them as ``utils.X`` everywhere.  We keep
that surface (the repo-root ``utils.py`` exposes the same names with the same
defaults) but internally carry an immutable-ish :class:`Settings` record so that
library code never depends on import-time globals.  Every field can also be
overridden from the environment as ``MPB_<FIELD>`` (e.g. ``MPB_MODEL_NAME=resnet34``),
which is what the launch scripts and tests use instead of editing a file.
"""
from __future__ import annotations

import dataclasses
import os
import types
from typing import Any, Dict

_TRUE = {"1", "true", "yes", "on"}


@dataclasses.dataclass
class Settings:
    # --- names shared with the reference (utils.py:4-45), same defaults ---------
    MODEL_NAME: str = "resnet18"
    FROM_CHECKPOINT: bool = False
    VALIDATE: bool = True
    DEBUG: bool = True
    N_IMAGES: int = 50000
    TRAIN_DIR: str = "./data/train/"
    TRAIN_FILE: str = "metadata.json"
    TEST_DIR: str = "./data/test/"
    CHECKPOINT_DIR: str = "./checkpoints/"
    MODELS_DIR: str = "./models/"
    WIDTH: int = 128
    HEIGHT: int = 128
    NUM_CLASSES: int = 64500
    BATCH_SIZE: int = 128
    LR: float = 4e-4
    NUM_EPOCHS: int = 10
    FEATURE_EXTRACT: bool = False
    USE_PRETRAINED: bool = True
    # --- new knobs (B200 runtime) ------------------------------------------------
    DEVICE: str = "auto"            # auto | cuda | cpu
    DTYPE: str = "bf16"             # bf16 | fp32 (compute dtype on CUDA)
    SYNTHETIC: bool = True          # on-device synthetic shards instead of JPEG files
    SYNTHETIC_SAMPLES: int = 0      # 0 -> derive from DEBUG (800) / full (40000) like main.py:73-82
    USE_INPUT_SIZE: bool = False    # honour the factory's input_size instead of WIDTH/HEIGHT
    OPTIMIZER: str = "adam"         # adam | sgd
    MOMENTUM: float = 0.9
    BUCKET_MB: float = 32.0         # gradient bucket size for the fused all-reduce
    ALLREDUCE_ALGO: str = "auto"    # auto | oneshot | twoshot | nvls | nccl
    GRAD_DTYPE: str = "fp32"        # fp32 | bf16 wire/accumulate dtype of gradient buckets
    NATIVE_OPS: bool = True         # run the hand-written sm_100a kernels on CUDA
    USE_CUDA_GRAPH: bool = True     # Trainer.fit replays the whole step as one CUDA graph for full batches
    SEED: int = 0
    TRAIN_CSV: str = "./data/train_sample.csv"
    TEST_CSV: str = "./data/test_sample.csv"
    LOG_DIR: str = "./logs/"
    PIPE_MICROBATCH: int = 32       # images per p2p message in the eval pipeline

    @property
    def CHECKPOINT_NAME(self) -> str:
        # utils.py:7 derives this from MODEL_NAME at import time; a property keeps it in sync.
        return "checkpoint_{}.pt".format(self.MODEL_NAME)

    def replace(self, **kw: Any) -> "Settings":
        return dataclasses.replace(self, **kw)

    def as_dict(self) -> Dict[str, Any]:
        d = dataclasses.asdict(self)
        d["CHECKPOINT_NAME"] = self.CHECKPOINT_NAME
        return d


def _coerce(value: str, like: Any) -> Any:
    if isinstance(like, bool):
        return value.strip().lower() in _TRUE
    if isinstance(like, int):
        return int(value)
    if isinstance(like, float):
        return float(value)
    return value


def from_module(mod=None, env: bool = True) -> Settings:
    """Build Settings from a constants module (the repo-root ``utils.py`` by default),
    then apply ``MPB_*`` environment overrides."""
    s = Settings()
    if mod is False:
        mod = None
    elif mod is None:
        try:
            import utils as mod  # type: ignore  # the reference-shaped constants module
        except Exception:  # pragma: no cover - no constants module on the path
            mod = None
    if mod is not None and hasattr(mod, "MODEL_NAME"):
        for f in dataclasses.fields(Settings):
            if hasattr(mod, f.name):
                setattr(s, f.name, getattr(mod, f.name))
    if env:
        for f in dataclasses.fields(Settings):
            key = "MPB_" + f.name
            if key in os.environ:
                setattr(s, f.name, _coerce(os.environ[key], getattr(s, f.name)))
    return s


def image_hw(settings: Settings, input_size: int | None = None) -> tuple[int, int]:
    """Height/width fed to the model.  The reference ignores the factory's ``input_size``
    (main.py:121 vs :64); we do the same unless USE_INPUT_SIZE is set."""
    if settings.USE_INPUT_SIZE and input_size:
        return int(input_size), int(input_size)
    return int(settings.HEIGHT), int(settings.WIDTH)
