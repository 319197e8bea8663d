"""Peer-memory collectives: the host runtime around ``csrc/allreduce.cu``.

One process per GPU.  Every arena handed out by :meth:`PeerCollectives.alloc` is a plain
``cudaMalloc`` allocation whose ``cudaIpcMemHandle`` is exchanged through the bootstrap
channel (``torch.distributed`` object all-gather -- plumbing only) and opened by every
peer, so each rank ends up with a table ``ptrs[rank] -> device pointer`` for every
symmetric buffer.  A view of such an arena is turned into "the same range on every rank"
by adding its byte offset to each table entry.  The kernels then load / store through
those pointers over NVLink 5 / NVSwitch and synchronise with system-scope release /
acquire flags in a symmetric signal pad (no NCCL on this path).

Algorithm choice per bucket (``auto``): one-shot up to ``ONESHOT_MAX_BYTES`` (latency
bound: one barrier pair, every rank reduces everything), two-shot above (bandwidth
bound: reduce-scatter by loads + all-gather by stores, optimizer state sliced ZeRO-1
style).  World size 1 runs the same kernels degenerate (W=1), which is what the single
GPU numerics tests exercise.
"""
from __future__ import annotations

import ctypes
import os
import socket
from typing import Dict, List, Optional, Tuple

import torch
import torch.distributed as dist

from ..build import lib_path
from .collectives import Hyper

ONESHOT_MAX_BYTES = 256 * 1024      # fallback crossover when no measured table covers (world, bytes)
ALGO_NAMES = {"oneshot": 0, "twoshot": 1, "nvls": 2}
ALGO_TABLE_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "algo_table.json")
c_void_p, c_int, c_ll, c_float = ctypes.c_void_p, ctypes.c_int, ctypes.c_longlong, ctypes.c_float


def load_algo_table(path: str = ALGO_TABLE_PATH) -> Dict[int, List[Tuple[int, str, int]]]:
    """The measured per-(world, bucket bytes) choice of algorithm and grid, written by ``scripts/test_peer.py --sweep``
    on the 8xB200 box: ``{"8": [[max_bytes, "oneshot|twoshot|nvls", blocks], ...]}`` with rows sorted by size; the
    first row whose ``max_bytes`` covers the bucket wins, the last row covers everything larger."""
    try:
        import json
        raw = json.load(open(path))
        return {int(w): [(int(b), str(a), int(g)) for b, a, g in rows] for w, rows in raw.get("table", {}).items()}
    except (OSError, ValueError, KeyError):
        return {}


def table_choice(table, world: int, nbytes: int, nvls: bool) -> Optional[Tuple[str, int]]:
    rows = table.get(int(world))
    if not rows:
        return None
    for max_bytes, algo, blocks in rows:
        if nbytes <= max_bytes or (max_bytes, algo, blocks) == rows[-1]:
            if algo == "nvls" and not nvls:
                algo = "twoshot"
            return algo, blocks
    return None


class _RawCuda:
    """Expose a raw device pointer through ``__cuda_array_interface__`` so torch can view it."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


class _SymBuf:
    def __init__(self, base: int, nbytes: int, peers: List[int], tensor: torch.Tensor, mc: int = 0):
        self.base, self.nbytes, self.peers, self.tensor, self.mc = base, nbytes, peers, tensor, mc


class _FdRelay:
    """Passes POSIX file descriptors (VMM / multicast handles) between the per-GPU processes.

    Rank 0 listens on an abstract AF_UNIX socket; every other rank keeps one connection to it.  An exchange
    sends each rank's descriptor to rank 0 (SCM_RIGHTS), which then hands every rank the full list."""

    _seq = 0                  # several relays may live in one process (trainer, then the eval pipeline's collectives)

    def __init__(self, world):
        self.rank, self.size = world.rank, world.size
        name = [None]
        _FdRelay._seq += 1
        if self.rank == 0:
            try:
                cand = "\0mpb-fd-{}-{}-{}".format(os.getpid(), os.environ.get("MASTER_PORT", "0"), _FdRelay._seq)
                self.srv = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
                self.srv.bind(cand)
                self.srv.listen(self.size)
                name[0] = cand
            except OSError:
                name[0] = None            # every rank learns of the failure below: nobody is left waiting
        dist.broadcast_object_list(name, src=0)
        if name[0] is None:
            raise OSError("fd relay: rank 0 could not open its listening socket")
        self.conns: Dict[int, socket.socket] = {}
        if self.rank == 0:
            for _ in range(self.size - 1):
                c, _addr = self.srv.accept()
                r = int.from_bytes(c.recv(4), "little")
                self.conns[r] = c
        else:
            c = socket.socket(socket.AF_UNIX, socket.SOCK_STREAM)
            c.connect(name[0])
            c.sendall(self.rank.to_bytes(4, "little"))
            self.conns[0] = c

    def all_gather_fd(self, fd: int) -> List[int]:
        """Every rank contributes one fd; returns the W descriptors (valid in this process) indexed by rank."""
        if self.rank == 0:
            fds = [fd] + [-1] * (self.size - 1)
            for r, c in self.conns.items():
                _msg, got, _flags, _addr = socket.recv_fds(c, 16, 1)
                fds[r] = got[0]
            for r, c in self.conns.items():
                socket.send_fds(c, [b"fds"], fds)
            return fds
        c = self.conns[0]
        socket.send_fds(c, [b"fd"], [fd])
        _msg, got, _flags, _addr = socket.recv_fds(c, 16, self.size)
        return list(got)

    def broadcast_fd(self, fd: int) -> int:
        if self.rank == 0:
            for c in self.conns.values():
                socket.send_fds(c, [b"mc"], [fd])
            return fd
        _msg, got, _flags, _addr = socket.recv_fds(self.conns[0], 16, 1)
        return got[0]


class PeerCollectives:
    name = "peer"

    def __init__(self, world, comm_blocks: int = 64):
        if world.device.type != "cuda":
            raise RuntimeError("PeerCollectives needs CUDA devices")
        self.world, self.size, self.rank, self.device = world, world.size, world.rank, world.device
        self.lib = ctypes.CDLL(lib_path())
        self.lib.mpb_oneshot_max_elems.restype = c_ll
        self.comm_blocks = comm_blocks
        self.algo_table = load_algo_table() if os.environ.get("MPB_ALGO_TABLE", "1") != "0" else {}
        self._ptr_cache: Dict[Tuple[int, int], object] = {}
        self.bufs: List[_SymBuf] = []
        self.launches = 0
        self.step_dev: Optional[torch.Tensor] = None    # device-resident optimizer step (set by the DP engine)
        torch.cuda.set_device(self.device)
        self.lib.mpb_vmm_round_size.restype = c_ll
        self.nvls = False
        self._relay: Optional[_FdRelay] = None
        if self.size > 1 and os.environ.get("MPB_NVLS", "1") != "0":
            ok = torch.tensor([int(self.lib.mpb_vmm_multicast_supported())], device=self.device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
            if int(ok.item()) == 1:
                try:
                    self._relay = _FdRelay(world)
                    self.nvls = True
                except OSError:
                    self.nvls = False
        pad_bytes = self.lib.mpb_signal_pad_bytes()
        self.pad = self._sym_alloc(pad_bytes)
        self.epochs = torch.zeros(self.lib.mpb_epoch_bytes() // 4, dtype=torch.int32, device=self.device)
        self._pads_arr = (c_void_p * self.size)(*self.pad.peers)
        self._count_stage = self._sym_alloc(256)
        if self.size > 1:
            dist.barrier()

    # ------------------------------------------------------------------ symmetric memory
    def _sym_alloc_vmm(self, nbytes: int) -> _SymBuf:
        """VMM allocation shared by fd + an NVLS multicast object bound over all ranks' copies."""
        size = int(self.lib.mpb_vmm_round_size(c_ll(nbytes), c_int(self.size)))
        ptr, fd, handle = c_void_p(), c_int(-1), ctypes.c_ulonglong(0)
        rc = self.lib.mpb_vmm_alloc(c_ll(size), ctypes.byref(ptr), ctypes.byref(fd), ctypes.byref(handle))
        if rc != 0:
            raise RuntimeError("mpb_vmm_alloc({}) failed: {}".format(size, rc))
        base = int(ptr.value)
        fds = self._relay.all_gather_fd(fd.value)
        peers = []
        for r in range(self.size):
            if r == self.rank:
                peers.append(base)
                continue
            p = c_void_p()
            rc = self.lib.mpb_vmm_import(c_int(fds[r]), c_ll(size), ctypes.byref(p))
            if rc != 0:
                raise RuntimeError("mpb_vmm_import(rank {}) failed: {}".format(r, rc))
            peers.append(int(p.value))
        for r, f in enumerate(fds):
            if f >= 0:                                   # every received descriptor, our own duplicate included
                try:
                    os.close(f)
                except OSError:
                    pass
        mc = ctypes.c_ulonglong(0)
        if self.rank == 0:
            mfd = c_int(-1)
            rc = self.lib.mpb_mc_create(c_ll(size), c_int(self.size), ctypes.byref(mfd), ctypes.byref(mc))
            if rc != 0:
                raise RuntimeError("cuMulticastCreate failed: {}".format(rc))
            self._relay.broadcast_fd(mfd.value)
            os.close(mfd.value)
        else:
            mfd = self._relay.broadcast_fd(-1)
            rc = self.lib.mpb_mc_import(c_int(mfd), ctypes.byref(mc))
            os.close(mfd)
            if rc != 0:
                raise RuntimeError("multicast import failed: {}".format(rc))
        rc = self.lib.mpb_mc_add_device(mc)
        if rc != 0:
            raise RuntimeError("cuMulticastAddDevice failed: {}".format(rc))
        dist.barrier()                                   # every device added before anyone binds
        mc_ptr = c_void_p()
        rc = self.lib.mpb_mc_bind_map(mc, handle, c_ll(size), ctypes.byref(mc_ptr))
        if rc != 0:
            raise RuntimeError("cuMulticastBindMem/map failed: {}".format(rc))
        dist.barrier()
        t = torch.as_tensor(_RawCuda(base, size), device=self.device)
        buf = _SymBuf(base, size, peers, t, int(mc_ptr.value))
        self.bufs.append(buf)
        return buf

    def _sym_alloc(self, nbytes: int) -> _SymBuf:
        if self.nvls:
            return self._sym_alloc_vmm(nbytes)
        nbytes = (int(nbytes) + 255) // 256 * 256
        ptr = c_void_p()
        rc = self.lib.mpb_sym_alloc(ctypes.byref(ptr), c_ll(nbytes))
        if rc != 0:
            raise RuntimeError("mpb_sym_alloc({}) failed: {}".format(nbytes, rc))
        base = int(ptr.value)
        peers = [base]
        if self.size > 1:
            hb = self.lib.mpb_ipc_handle_bytes()
            hbuf = ctypes.create_string_buffer(hb)
            rc = self.lib.mpb_ipc_get_handle(c_void_p(base), hbuf)
            if rc != 0:
                raise RuntimeError("cudaIpcGetMemHandle failed: {}".format(rc))
            handles: List[Optional[bytes]] = [None] * self.size
            dist.all_gather_object(handles, bytes(hbuf.raw))
            peers = []
            for r, h in enumerate(handles):
                if r == self.rank:
                    peers.append(base)
                    continue
                p = c_void_p()
                rc = self.lib.mpb_ipc_open_handle(ctypes.create_string_buffer(h, hb), ctypes.byref(p))
                if rc != 0:
                    raise RuntimeError("cudaIpcOpenMemHandle(rank {}) failed: {}".format(r, rc))
                peers.append(int(p.value))
        t = torch.as_tensor(_RawCuda(base, nbytes), device=self.device)
        buf = _SymBuf(base, nbytes, peers, t)
        self.bufs.append(buf)
        return buf

    def alloc(self, numel: int, dtype: torch.dtype) -> torch.Tensor:
        esz = torch.empty((), dtype=dtype).element_size()
        buf = self._sym_alloc(numel * esz)
        return buf.tensor[: numel * esz].view(dtype)

    def _peer_ptrs(self, view: Optional[torch.Tensor]):
        """ctypes array of W pointers addressing ``view``'s range on every rank (or None)."""
        if view is None:
            return None
        p = view.data_ptr()
        hit = self._ptr_cache.get((p, 0))         # bucket views recur every step: build each pointer table once
        if hit is not None:
            return hit
        for b in self.bufs:
            if b.base <= p < b.base + b.nbytes:
                off = p - b.base
                arr = (c_void_p * self.size)(*[q + off for q in b.peers])
                self._ptr_cache[(p, 0)] = arr
                return arr
        if self.size == 1:
            return (c_void_p * 1)(p)
        raise ValueError("tensor is not inside a symmetric arena obtained from PeerCollectives.alloc")

    def _mc_ptr(self, view: Optional[torch.Tensor]) -> int:
        if view is None:
            return 0
        p = view.data_ptr()
        hit = self._ptr_cache.get((p, 1))
        if hit is not None:
            return hit
        for b in self.bufs:
            if b.mc and b.base <= p < b.base + b.nbytes:
                self._ptr_cache[(p, 1)] = b.mc + (p - b.base)
                return b.mc + (p - b.base)
        return 0

    # ------------------------------------------------------------------ collectives
    def _plan(self, algo: str, nbytes: int, numel: int) -> Tuple[int, int]:
        """(algorithm id, grid) for one bucket.  ``algo`` may be a name, ``name:blocks`` (sweeps), or ``auto`` = the
        measured table for this world size, falling back to the fixed crossover."""
        blocks = 0
        if ":" in algo:
            algo, b = algo.split(":", 1)
            blocks = int(b)
        if algo in ALGO_NAMES:
            a = ALGO_NAMES[algo]
            if a == 2 and not self.nvls:
                a = 1
        else:
            pick = table_choice(self.algo_table, self.size, nbytes, self.nvls)
            if pick is not None:
                a, blocks = ALGO_NAMES[pick[0]], (blocks or pick[1])
            elif self.nvls and nbytes > ONESHOT_MAX_BYTES:
                a = 2
            else:
                a = 0 if nbytes <= ONESHOT_MAX_BYTES else 1
        if a == 0 and numel > self.lib.mpb_oneshot_max_elems(c_int(blocks or self._blocks(numel, 0))):
            a = 1
        if not blocks:
            blocks = self._blocks(numel, a)
        elif a != 0:
            blocks = max(1, min(blocks, (numel // max(self.size, 1) + 2047) // 2048))
        return a, int(blocks)

    def _algo(self, algo: str, nbytes: int, numel: int) -> int:
        return self._plan(algo, nbytes, numel)[0]

    def _blocks(self, numel: int, algo: int) -> int:
        per = 512 * 4
        work = numel if algo == 0 else numel // max(self.size, 1)
        need = max(1, (work + per - 1) // per)
        cap = 128 if (algo == 0 or self.size == 1) else self.comm_blocks
        return int(min(cap, need))

    def owns_full_state(self, algo: str) -> bool:
        # conservative: any bucket that may run two-shot keeps sliced optimizer state
        return self.size == 1

    def _launch_ar(self, grads, params, shadows, m, v, hyper: Optional[Hyper], step: int, algo: str,
                   wire: Optional[torch.Tensor] = None) -> None:
        n = grads.numel()
        a, blocks = self._plan(algo, n * (2 if wire is not None else 4), n)
        p = self._peer_ptrs(params)
        s = self._peer_ptrs(shadows)
        kind = 0 if hyper is None else (1 if hyper.kind == "adam" else 2)
        h = hyper or Hyper()
        step_dev = c_void_p(0 if self.step_dev is None else self.step_dev.data_ptr())
        stream = c_void_p(torch.cuda.current_stream().cuda_stream)
        if wire is not None:
            # bf16 wire: the kernel packs this rank's fp32 bucket into `wire`, peers read 2 bytes per element
            if kind == 0:
                raise ValueError("the bf16 wire format is only defined for the fused optimizer epilogue")
            a = 2 if a == 2 else 1
            mc_w = mc_p = mc_s = 0
            if a == 2:
                mc_w, mc_p, mc_s = self._mc_ptr(wire), self._mc_ptr(params), self._mc_ptr(shadows)
                if mc_w == 0 or mc_p == 0 or (shadows is not None and mc_s == 0):
                    a = 1
            rc = self.lib.mpb_allreduce_fused_bf16wire(
                c_void_p(grads.data_ptr()), self._peer_ptrs(wire), p, s, self._pads_arr,
                c_void_p(self.epochs.data_ptr()), c_void_p(m.data_ptr()), c_void_p(0 if v is None else v.data_ptr()),
                c_ll(n), c_int(self.rank), c_int(self.size), c_int(a), c_int(blocks), c_int(kind), c_int(step),
                c_float(h.lr), c_float(h.beta1), c_float(h.beta2), c_float(h.eps), c_float(h.weight_decay),
                c_float(h.momentum), step_dev, c_void_p(mc_w), c_void_p(mc_p), c_void_p(mc_s), stream)
            self.launches += 1
            if rc != 0:
                raise RuntimeError("mpb_allreduce_fused_bf16wire failed: {}".format(rc))
            return
        g = self._peer_ptrs(grads)
        mc_g = mc_p = mc_s = 0
        if a == 2:
            mc_g, mc_p, mc_s = self._mc_ptr(grads), self._mc_ptr(params), self._mc_ptr(shadows)
            if mc_g == 0 or (kind != 0 and mc_p == 0) or (shadows is not None and kind != 0 and mc_s == 0):
                a = 1                                     # range not covered by a multicast mapping
        rc = self.lib.mpb_allreduce_fused(
            g, p, s, self._pads_arr, c_void_p(self.epochs.data_ptr()),
            c_void_p(0 if m is None else m.data_ptr()), c_void_p(0 if v is None else v.data_ptr()),
            c_ll(n), c_int(self.rank), c_int(self.size), c_int(a), c_int(blocks), c_int(kind), c_int(step),
            c_float(h.lr), c_float(h.beta1), c_float(h.beta2), c_float(h.eps), c_float(h.weight_decay),
            c_float(h.momentum), step_dev, c_void_p(mc_g), c_void_p(mc_p), c_void_p(mc_s), stream)
        self.launches += 1
        if rc != 0:
            raise RuntimeError("mpb_allreduce_fused failed: {}".format(rc))

    def all_reduce_avg_(self, view: torch.Tensor, algo: str = "auto") -> None:
        if view.dtype == torch.bfloat16:
            return self._all_reduce_avg_bf16(view)
        if view.dtype != torch.float32:
            raise TypeError("peer all-reduce operates on fp32 or bf16 arenas")
        self._launch_ar(view, None, None, None, None, None, 1, algo)

    def _all_reduce_avg_bf16(self, view: torch.Tensor) -> None:
        """bf16 wire format: peers exchange 2 bytes per gradient element (fp32 accumulation inside the kernel)."""
        n = view.numel()
        if n % (8 * self.size) != 0:
            raise ValueError("bf16 all-reduce range must be a multiple of 8 * world elements")
        blocks = self._blocks(n // 2, 1)
        rc = self.lib.mpb_allreduce_bf16(self._peer_ptrs(view), self._pads_arr, c_void_p(self.epochs.data_ptr()), c_ll(n),
                                         c_int(self.rank), c_int(self.size), c_int(blocks),
                                         c_void_p(torch.cuda.current_stream().cuda_stream))
        self.launches += 1
        if rc != 0:
            raise RuntimeError("mpb_allreduce_bf16 failed: {}".format(rc))

    def all_reduce_update_(self, grads, params, state: Dict[str, torch.Tensor], shadow, hyper: Hyper, step: int,
                           algo: str = "auto", wire: Optional[torch.Tensor] = None) -> None:
        if hyper.kind == "adam":
            m, v = state["exp_avg"], state["exp_avg_sq"]
        else:
            m, v = state["momentum_buffer"], None
        self._launch_ar(grads, params, shadow, m, v, hyper, step, algo, wire)

    def broadcast_(self, view: torch.Tensor, root: int = 0) -> None:
        if self.size == 1:
            return
        nbytes = view.numel() * view.element_size()
        if nbytes % 16 != 0:
            raise ValueError("broadcast range must be a multiple of 16 bytes")
        n_floats = nbytes // 4
        rc = self.lib.mpb_broadcast(self._peer_ptrs(view), self._pads_arr, c_void_p(self.epochs.data_ptr()),
                                    c_ll(n_floats), c_int(self.rank), c_int(self.size), c_int(root),
                                    c_int(self._blocks(n_floats * self.size, 1)),
                                    c_void_p(torch.cuda.current_stream().cuda_stream))
        self.launches += 1
        if rc != 0:
            raise RuntimeError("mpb_broadcast failed: {}".format(rc))

    def all_gather_slices_(self, view: torch.Tensor) -> None:
        if self.size == 1:
            return
        # optimizer moments are ordinary (non-symmetric) tensors: stage through a symmetric scratch
        n = view.numel()
        if not hasattr(self, "_ag_stage") or self._ag_stage.numel() < n:
            self._ag_stage = self.alloc(n, torch.float32)
        stage = self._ag_stage[:n]
        stage.copy_(view)
        rc = self.lib.mpb_allgather_slices(self._peer_ptrs(stage), self._pads_arr, c_void_p(self.epochs.data_ptr()),
                                           c_ll(n), c_int(self.rank), c_int(self.size), c_int(self._blocks(n, 1)),
                                           c_void_p(torch.cuda.current_stream().cuda_stream))
        self.launches += 1
        if rc != 0:
            raise RuntimeError("mpb_allgather_slices failed: {}".format(rc))
        view.copy_(stage)

    def all_reduce_sum_i64(self, value: torch.Tensor) -> torch.Tensor:
        stage = self._count_stage.tensor[:8].view(torch.int64)
        stage.copy_(value.reshape(1))
        out = torch.empty(1, dtype=torch.int64, device=self.device)
        vals = (c_void_p * self.size)(*self._count_stage.peers)
        rc = self.lib.mpb_count_allreduce(vals, self._pads_arr, c_void_p(self.epochs.data_ptr()),
                                          c_void_p(out.data_ptr()), c_int(self.rank), c_int(self.size),
                                          c_void_p(torch.cuda.current_stream().cuda_stream))
        self.launches += 1
        if rc != 0:
            raise RuntimeError("mpb_count_allreduce failed: {}".format(rc))
        value.copy_(out.reshape(value.shape))
        return value

    def barrier(self) -> None:
        torch.cuda.synchronize(self.device)
        if self.size > 1:
            dist.barrier()

    def sync_stream(self) -> None:
        torch.cuda.current_stream(self.device).synchronize()

    def close(self) -> None:
        # allocations live for the life of the process (peers may still hold mappings); nothing to do
        pass
