"""Per-rank logger.

Same logger name and line prefix as the reference (``INFO:Herbarium_R{rank}:...``,
/root/reference/main.py:22-41, evaluation_pipeline.py:19-38) but created on demand
instead of at import time, and each rank appends to its *own* file
(``training.r{rank}.log``) so lines never interleave.  Rank 0 also keeps
the reference's un-suffixed file name so existing tooling that tails ``training.log``
still works.
"""
from __future__ import annotations

import logging
import os

_NAME = "Herbarium"


def init_logger(log_file: str = "training.log", rank: int = 0, log_dir: str | None = None,
                stream: bool = True) -> logging.Logger:
    logger = logging.getLogger(_NAME)
    logger.setLevel(logging.DEBUG)
    for h in list(logger.handlers):          # idempotent re-initialisation
        logger.removeHandler(h)
        try:
            h.close()
        except Exception:
            pass
    fmt = logging.Formatter("%(levelname)s:%(name)s_R{}:%(message)s".format(rank))
    if stream:
        sh = logging.StreamHandler()
        sh.setLevel(logging.DEBUG)
        sh.setFormatter(fmt)
        logger.addHandler(sh)
    if log_file:
        base, ext = os.path.splitext(log_file)
        if log_dir:
            os.makedirs(log_dir, exist_ok=True)
            base = os.path.join(log_dir, os.path.basename(base))
        path = (base + ext) if rank == 0 else "{}.r{}{}".format(base, rank, ext)
        try:
            fh = logging.FileHandler(path)
            fh.setFormatter(fmt)
            logger.addHandler(fh)
        except OSError:  # read-only working directory: keep the stream handler only
            pass
    logger.propagate = False
    return logger


def get_logger() -> logging.Logger:
    logger = logging.getLogger(_NAME)
    if not logger.handlers:
        logger.addHandler(logging.NullHandler())
    return logger
