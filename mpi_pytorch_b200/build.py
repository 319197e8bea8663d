"""In-tree build of the native library ``mpi_pytorch_b200/lib/libmpb200.so``.

Every ``csrc/*.cu`` is compiled with ``nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo``
(sm_100a only -- there is no multi-arch fat binary and no fallback path) into one shared object
with a plain C ABI (``include/mpb200.h``), loaded through ``ctypes`` by
``mpi_pytorch_b200/ops/cuda_impl.py`` and ``parallel/peer.py``.  nvcc cross-compiles without a GPU, so
``build()`` runs on the CPU-only build box; the resulting ``.so`` travels to the GPU box in-tree.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys
from typing import List

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
OBJDIR = os.path.join(LIBDIR, "obj")
LIB = os.path.join(LIBDIR, "libmpb200.so")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-lineinfo", "-std=c++17", "--use_fast_math", "-Xcompiler", "-fPIC", "-Xcompiler",
              "-fvisibility=hidden", "--expt-relaxed-constexpr", "-Xptxas", "-v"]


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def _sources() -> List[str]:
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest(paths: List[str]) -> str:
    h = hashlib.sha256()
    for p in paths:
        with open(p, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS + ARCH).encode())
    return h.hexdigest()


def _compile_one(nvcc: str, src: str, headers: List[str], verbose: bool) -> str:
    obj = os.path.join(OBJDIR, os.path.basename(src)[:-3] + ".o")
    stamp = obj + ".sha"
    dig = _digest([src] + headers)
    if os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dig:
        return obj
    cmd = [nvcc] + ARCH + NVCC_FLAGS + ["-I", CSRC, "-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    log = os.path.join(OBJDIR, os.path.basename(src)[:-3] + ".log")
    with open(log, "w") as fh:
        fh.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for {}:\n{}".format(src, r.stderr[-4000:]))
    if verbose:
        print("[mpb build] compiled", os.path.basename(src))
    with open(stamp, "w") as fh:
        fh.write(dig)
    return obj


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJDIR, exist_ok=True)
    nvcc = nvcc_path()
    srcs = _sources()
    headers = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h")))
    if force:
        for f in os.listdir(OBJDIR):
            os.remove(os.path.join(OBJDIR, f))
    with cf.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        objs = list(ex.map(lambda s: _compile_one(nvcc, s, headers, verbose), srcs))
    link_stamp = LIB + ".sha"
    dig = _digest(objs)
    if not (os.path.exists(LIB) and os.path.exists(link_stamp) and open(link_stamp).read() == dig):
        cmd = [nvcc] + ARCH + ["-shared", "-o", LIB] + objs      # libcuda entry points are resolved at run time
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stderr[-4000:])
        with open(link_stamp, "w") as fh:
            fh.write(dig)
        if verbose:
            print("[mpb build] linked", LIB)
    return LIB


def lib_path(build_if_missing: bool = True) -> str:
    if not os.path.exists(LIB):
        if not build_if_missing:
            raise FileNotFoundError(LIB)
        build()
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
