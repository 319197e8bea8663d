"""Minimal stand-in for the ``mpi4py`` package, used ONLY by the reference arm of bench.py.

Neither the build box nor the GPU box has mpi4py, ``mpiexec`` or any MPI library, so the
unmodified reference (``baseline/_ref``: ``from mpi4py import MPI``) could not even be imported.
This shim supplies the handful of ``MPI.COMM_WORLD`` calls the reference makes
(/root/reference/mpi_tools.py, main.py, evaluation_pipeline.py) on top of ``torch.distributed``
with the **gloo** (CPU, host-memory) backend -- the closest available equivalent of the host MPI
transport the reference was written for.  Rank/size come from the torchrun environment.
It is a substitute for a missing third-party dependency, not a change to the reference.
"""
import os
import pickle
import time

import numpy as np
import torch
import torch.distributed as dist


class _Op:
    def __init__(self, name):
        self.name = name


class _Comm:
    def _ensure(self):
        size = self.Get_size()
        if size > 1 and not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("MASTER_PORT", "29511")
            dist.init_process_group("gloo", rank=self.Get_rank(), world_size=size)
        return size

    # rank/size as a launcher would export them: torchrun's names first, then Open MPI's (what ``mpiexec`` sets)
    def Get_rank(self):
        return int(os.environ.get("RANK", os.environ.get("OMPI_COMM_WORLD_RANK", "0")))

    def Get_size(self):
        return int(os.environ.get("WORLD_SIZE", os.environ.get("OMPI_COMM_WORLD_SIZE", "1")))

    # ---- buffer collectives (numpy arrays, in the style of mpi4py's upper-case methods)
    def Allreduce(self, sendbuf, recvbuf, op=None):
        if self._ensure() == 1:
            np.copyto(recvbuf, sendbuf)
            return
        t = torch.from_numpy(np.ascontiguousarray(sendbuf).copy())
        dist.all_reduce(t)
        np.copyto(recvbuf, t.numpy().reshape(recvbuf.shape))

    def Bcast(self, buf, root=0):
        if self._ensure() == 1:
            return
        t = torch.from_numpy(buf) if buf.flags["C_CONTIGUOUS"] else torch.from_numpy(np.ascontiguousarray(buf))
        dist.broadcast(t, src=root)
        if not buf.flags["C_CONTIGUOUS"]:
            np.copyto(buf, t.numpy())

    # ---- pickled object collectives (lower-case methods)
    def scatter(self, objs, root=0):
        if self._ensure() == 1:
            return objs[0]
        out = [None]
        dist.scatter_object_list(out, objs if self.Get_rank() == root else None, src=root)
        return out[0]

    def bcast(self, obj, root=0):
        if self._ensure() == 1:
            return obj
        box = [obj]
        dist.broadcast_object_list(box, src=root)
        return box[0]

    def send(self, obj, dest):
        self._ensure()
        payload = torch.frombuffer(bytearray(pickle.dumps(obj)), dtype=torch.uint8)
        dist.send(torch.tensor([payload.numel()], dtype=torch.int64), dst=dest)
        dist.send(payload, dst=dest)

    def recv(self, source=0):
        self._ensure()
        n = torch.zeros(1, dtype=torch.int64)
        dist.recv(n, src=source)
        payload = torch.empty(int(n.item()), dtype=torch.uint8)
        dist.recv(payload, src=source)
        return pickle.loads(payload.numpy().tobytes())

    def reduce(self, value, op=None, root=0):
        if self._ensure() == 1:
            return value
        vals = [None] * self.Get_size() if self.Get_rank() == root else None
        dist.gather_object(value, vals, dst=root)
        if vals is None:
            return None
        total = vals[0]
        for v in vals[1:]:
            total = total + v
        return total

    def Barrier(self):
        if self._ensure() > 1:
            dist.barrier()


class _MPI:
    SUM = _Op("sum")
    COMM_WORLD = _Comm()

    @staticmethod
    def Wtime():
        return time.perf_counter()


MPI = _MPI()
