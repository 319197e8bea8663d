"""Rank bootstrap: one process per GPU.

The reference obtains ``rank``/``size`` from ``MPI.COMM_WORLD`` at import time
(/root/reference/main.py:16-18, evaluation_pipeline.py:13-15, mpi_tools.py:5-9).
mpi4py is kept *only* for that purpose here: if it is importable and the process was
started by ``mpiexec`` we read rank/size from it; otherwise we read the launcher's
environment (torchrun ``RANK/WORLD_SIZE/LOCAL_RANK``, Open MPI ``OMPI_COMM_WORLD_*``,
PMI ``PMI_RANK/PMI_SIZE``, Slurm) and finally fall back to a single process or a
self-spawned local job (:func:`spawn`).  All data movement then goes through
``torch.distributed`` (NCCL on GPUs, gloo on CPU) and our own peer-memory kernels.

:class:`Comm` is an mpi4py-flavoured facade (``Get_rank``, ``scatter``, ``send`` ...)
so code written against ``comm = MPI.COMM_WORLD`` keeps its shape.
"""
from __future__ import annotations

import dataclasses
import datetime
import os
import socket
import time
from typing import Any, Callable, List, Optional, Sequence

import torch
import torch.distributed as dist

_ENV_RANK = ("RANK", "OMPI_COMM_WORLD_RANK", "PMI_RANK", "PMIX_RANK", "SLURM_PROCID")
_ENV_SIZE = ("WORLD_SIZE", "OMPI_COMM_WORLD_SIZE", "PMI_SIZE", "SLURM_NTASKS")
_ENV_LOCAL = ("LOCAL_RANK", "OMPI_COMM_WORLD_LOCAL_RANK", "MPI_LOCALRANKID", "SLURM_LOCALID")


@dataclasses.dataclass
class World:
    rank: int = 0
    size: int = 1
    local_rank: int = 0
    device: torch.device = dataclasses.field(default_factory=lambda: torch.device("cpu"))
    backend: Optional[str] = None     # nccl | gloo | None (single process)
    source: str = "single"            # mpi4py | env | single | spawn

    @property
    def is_root(self) -> bool:
        return self.rank == 0


_WORLD: Optional[World] = None


def _first_env(names: Sequence[str]) -> Optional[int]:
    for n in names:
        v = os.environ.get(n)
        if v is not None and v != "":
            try:
                return int(v)
            except ValueError:
                continue
    return None


def _probe_mpi() -> Optional[tuple]:
    """Return (rank, size, comm) from mpi4py when it exists AND we run under an MPI launcher."""
    if not any(k in os.environ for k in ("OMPI_COMM_WORLD_SIZE", "PMI_SIZE", "PMIX_RANK", "MPI_LOCALRANKID")):
        return None
    try:
        from mpi4py import MPI  # type: ignore
    except Exception:
        return None
    comm = MPI.COMM_WORLD
    return comm.Get_rank(), comm.Get_size(), comm


def pick_device(kind: str, local_rank: int) -> torch.device:
    if kind == "cpu":
        return torch.device("cpu")
    if kind in ("auto", "cuda") and torch.cuda.is_available():
        n = torch.cuda.device_count()
        dev = torch.device("cuda", local_rank % max(n, 1))
        torch.cuda.set_device(dev)
        return dev
    if kind == "cuda":
        raise RuntimeError("DEVICE='cuda' requested but no CUDA device is visible")
    return torch.device("cpu")


def free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return int(s.getsockname()[1])


def init_world(device: str = "auto", timeout_s: int = 1800) -> World:
    """Idempotent.  Discovers rank/size, binds the device, initialises torch.distributed."""
    global _WORLD
    if _WORLD is not None:
        return _WORLD
    source = "single"
    rank, size, local = 0, 1, 0
    mpi = _probe_mpi()
    if mpi is not None:
        rank, size, comm = mpi
        local = _first_env(_ENV_LOCAL) or 0
        source = "mpi4py"
        if size > 1 and "MASTER_ADDR" not in os.environ:
            # rank 0 picks the rendezvous endpoint, everyone learns it over MPI (bootstrap only)
            ep = ("127.0.0.1", free_port()) if rank == 0 else None
            ep = comm.bcast(ep, root=0)
            os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = ep[0], str(ep[1])
    else:
        r, s = _first_env(_ENV_RANK), _first_env(_ENV_SIZE)
        if r is not None and s is not None:
            rank, size = r, s
            local = _first_env(_ENV_LOCAL)
            if local is None:
                local = rank
            source = "env"
    dev = pick_device(device, local)
    backend = None
    if size > 1:
        backend = "nccl" if dev.type == "cuda" else "gloo"
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29500")
            kw = dict(backend=backend, rank=rank, world_size=size,
                      timeout=datetime.timedelta(seconds=timeout_s))
            if backend == "nccl":
                kw["device_id"] = dev
            dist.init_process_group(**kw)
    _WORLD = World(rank=rank, size=size, local_rank=local, device=dev, backend=backend, source=source)
    return _WORLD


def world() -> World:
    return _WORLD if _WORLD is not None else init_world()


def shutdown() -> None:
    global _WORLD
    if dist.is_available() and dist.is_initialized():
        try:
            dist.barrier()
        except Exception:
            pass
        dist.destroy_process_group()
    _WORLD = None


def _spawn_entry(local_rank: int, nprocs: int, port: int, device: str, fn: Callable, args: tuple):
    os.environ.update({"RANK": str(local_rank), "WORLD_SIZE": str(nprocs), "LOCAL_RANK": str(local_rank),
                       "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port)})
    torch.set_num_threads(1)
    w = init_world(device)
    w.source = "spawn"
    try:
        fn(w, *args)
    finally:
        shutdown()


def spawn(fn: Callable, nprocs: int, args: tuple = (), device: str = "auto") -> None:
    """Self-spawn ``nprocs`` local ranks (the no-MPI replacement for ``mpiexec -n N``).
    ``fn(world, *args)`` runs in every rank."""
    import torch.multiprocessing as mp
    if nprocs <= 1:
        w = init_world(device)
        fn(w, *args)
        return
    mp.spawn(_spawn_entry, args=(nprocs, free_port(), device, fn, args), nprocs=nprocs, join=True)


class Comm:
    """mpi4py-shaped communicator over torch.distributed.

    Lower-case methods move picklable Python objects (like mpi4py's), upper-case ones
    move tensors / buffers in place.  Object p2p (`send`/`recv`) is used only by the
    CPU fallback of the evaluation pipeline and by tests; the GPU pipeline sends
    fixed-shape device tensors (parallel/pipeline.py).
    """

    def __init__(self, w: Optional[World] = None):
        self.w = w or world()

    # -- identity ---------------------------------------------------------------
    def Get_rank(self) -> int:
        return self.w.rank

    def Get_size(self) -> int:
        return self.w.size

    @staticmethod
    def Wtime() -> float:
        return time.perf_counter()

    def Barrier(self) -> None:
        if self.w.size > 1:
            dist.barrier()

    # -- object collectives -----------------------------------------------------
    def scatter(self, objs: Optional[List[Any]], root: int = 0) -> Any:
        if self.w.size == 1:
            return objs[0] if objs is not None else None
        out: List[Any] = [None]
        dist.scatter_object_list(out, objs if self.w.rank == root else None, src=root)
        return out[0]

    def bcast(self, obj: Any, root: int = 0) -> Any:
        if self.w.size == 1:
            return obj
        box = [obj if self.w.rank == root else None]
        dist.broadcast_object_list(box, src=root)
        return box[0]

    def gather(self, obj: Any, root: int = 0) -> Optional[List[Any]]:
        if self.w.size == 1:
            return [obj]
        out = [None] * self.w.size if self.w.rank == root else None
        dist.gather_object(obj, out, dst=root)
        return out

    def reduce(self, value: Any, op: str = "sum", root: int = 0) -> Any:
        vals = self.gather(value, root=root)
        if vals is None:
            return None
        if op == "sum":
            total = vals[0]
            for v in vals[1:]:
                total = total + v
            return total
        if op == "max":
            return max(vals)
        raise ValueError(op)

    def send(self, obj: Any, dest: int) -> None:
        import io
        import pickle
        buf = io.BytesIO()
        pickle.dump(obj, buf)
        payload = torch.frombuffer(bytearray(buf.getvalue()), dtype=torch.uint8)
        n = torch.tensor([payload.numel()], dtype=torch.int64)
        if self.w.backend == "nccl":
            n, payload = n.to(self.w.device), payload.to(self.w.device)
        dist.send(n, dst=dest)
        dist.send(payload, dst=dest)

    def recv(self, source: int) -> Any:
        import io
        import pickle
        dev = self.w.device if self.w.backend == "nccl" else torch.device("cpu")
        n = torch.zeros(1, dtype=torch.int64, device=dev)
        dist.recv(n, src=source)
        payload = torch.empty(int(n.item()), dtype=torch.uint8, device=dev)
        dist.recv(payload, src=source)
        return pickle.load(io.BytesIO(payload.cpu().numpy().tobytes()))

    # -- buffer collectives -----------------------------------------------------
    def Allreduce(self, t: torch.Tensor) -> torch.Tensor:
        if self.w.size > 1:
            dist.all_reduce(t)
        return t

    def Bcast(self, t: torch.Tensor, root: int = 0) -> torch.Tensor:
        if self.w.size > 1:
            dist.broadcast(t, src=root)
        return t
