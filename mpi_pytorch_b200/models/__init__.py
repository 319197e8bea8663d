"""Model zoo (see :mod:`.factory`)."""
from .factory import MODEL_NAMES, initialize_model, set_parameter_requires_grad  # noqa: F401
