"""mpi_pytorch_b200: B200-native data-parallel CNN training and pipelined inference.

A from-scratch framework with the capabilities of erick093/MPI_Pytorch:
``models`` (zoo + factory), ``ops`` (sm_100a kernels + torch oracles), ``parallel``
(bootstrap, fused peer-memory all-reduce, DP engine, eval pipeline), ``data``,
``engine`` (trainer, checkpoints) and ``utils`` (config, logging, timing).
"""
__version__ = "0.1.0"
