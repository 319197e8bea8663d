"""Run a function on N gloo ranks (CPU) and collect per-rank results."""
import os
import pickle
import tempfile
import traceback

import torch
import torch.multiprocessing as mp


def _entry(rank, nprocs, port, fn, args, outdir):
    os.environ.update({"RANK": str(rank), "WORLD_SIZE": str(nprocs), "LOCAL_RANK": str(rank),
                       "MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "MPB_LOG_DIR": outdir})
    torch.set_num_threads(1)
    from mpi_pytorch_b200.parallel import bootstrap
    try:
        w = bootstrap.init_world("cpu")
        res = fn(w, *args)
        with open(os.path.join(outdir, "r%d.pkl" % rank), "wb") as fh:
            pickle.dump(res, fh)
    except Exception:
        traceback.print_exc()
        raise
    finally:
        bootstrap.shutdown()


def run_ranks(fn, nprocs, args=()):
    from mpi_pytorch_b200.parallel.bootstrap import free_port
    with tempfile.TemporaryDirectory() as d:
        mp.spawn(_entry, args=(nprocs, free_port(), fn, args, d), nprocs=nprocs, join=True)
        out = []
        for r in range(nprocs):
            with open(os.path.join(d, "r%d.pkl" % r), "rb") as fh:
                out.append(pickle.load(fh))
        return out
