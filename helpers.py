"""Checkpoint helpers front (reference surface: /root/reference/helpers.py:4-15)."""
from mpi_pytorch_b200.engine.checkpoint import (load_checkpoint, make_checkpoint,  # noqa: F401
                                                save_checkpoint)
