"""Config, logging and timing helpers."""
from .config import Settings, from_module, image_hw  # noqa: F401
from .logging import get_logger, init_logger  # noqa: F401
from .timing import ClockSampler, DeviceTimer, max_over_ranks, wtime  # noqa: F401
