"""Communication helpers front (reference surface: /root/reference/mpi_tools.py:5-53)."""
from mpi_pytorch_b200.parallel.mpi_tools import (SUM, mpi_all_reduce, mpi_avg_grads,  # noqa: F401
                                                 mpi_broadcast, mpi_sum, num_processes, sync_params)
