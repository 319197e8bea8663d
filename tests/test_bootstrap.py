"""Rank bootstrap: the ``mpiexec -n N python -m mpi4py main.py`` launch form (README.md:38 of the reference).

No MPI exists on the build or GPU boxes, so the mpi4py branch of ``bootstrap.init_world`` is driven here through
``baseline/mpi4py_shim`` (an ``MPI.COMM_WORLD`` over gloo) with the environment Open MPI's ``mpiexec`` would export."""
import os
import sys

import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _mpi_rank(rank, size, port, outdir):
    for k in ("RANK", "WORLD_SIZE", "LOCAL_RANK", "MASTER_ADDR"):
        os.environ.pop(k, None)
    os.environ.update({"OMPI_COMM_WORLD_RANK": str(rank), "OMPI_COMM_WORLD_SIZE": str(size),
                       "OMPI_COMM_WORLD_LOCAL_RANK": str(rank), "MASTER_PORT": str(port), "MPB_LOG_DIR": outdir})
    sys.path.insert(0, os.path.join(ROOT, "baseline", "mpi4py_shim"))
    torch.set_num_threads(1)
    from mpi_pytorch_b200.parallel import bootstrap, mpi_tools
    probe = bootstrap._probe_mpi()
    assert probe is not None and probe[:2] == (rank, size)
    w = bootstrap.init_world("cpu")
    assert (w.source, w.rank, w.size, w.local_rank, w.backend) == ("mpi4py", rank, size, rank, "gloo")
    assert "MASTER_ADDR" in os.environ                      # rank 0's endpoint arrived through comm.bcast
    assert mpi_tools.num_processes() == size
    assert mpi_tools.mpi_sum(float(rank + 1)) == float(sum(range(1, size + 1)))
    bootstrap.shutdown()
    open(os.path.join(outdir, "ok%d" % rank), "w").close()


def test_mpi4py_launch_path_through_shim(tmp_path):
    from mpi_pytorch_b200.parallel.bootstrap import free_port
    mp.spawn(_mpi_rank, args=(2, free_port(), str(tmp_path)), nprocs=2, join=True)
    assert os.path.exists(tmp_path / "ok0") and os.path.exists(tmp_path / "ok1")


def test_probe_ignores_mpi4py_without_a_launcher(monkeypatch):
    from mpi_pytorch_b200.parallel import bootstrap
    for k in ("OMPI_COMM_WORLD_SIZE", "PMI_SIZE", "PMIX_RANK", "MPI_LOCALRANKID"):
        monkeypatch.delenv(k, raising=False)
    assert bootstrap._probe_mpi() is None
