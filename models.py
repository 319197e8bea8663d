"""``models.initialize_model`` front (reference surface: /root/reference/models.py:16)."""
from mpi_pytorch_b200.models.factory import (MODEL_NAMES, initialize_model,  # noqa: F401
                                             set_parameter_requires_grad)
