"""Training engine and checkpoint I/O."""
from .checkpoint import load_checkpoint, make_checkpoint, save_checkpoint  # noqa: F401
