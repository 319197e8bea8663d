"""Dataset front (reference surface: /root/reference/data_loader.py:6)."""
from mpi_pytorch_b200.data.dataset import GetData  # noqa: F401
