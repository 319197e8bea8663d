"""Data-parallel training entry point (reference surface: /root/reference/main.py).

Launch forms:
  mpiexec -n N python -m mpi4py main.py        # where MPI + mpi4py exist (bootstrap only)
  torchrun --nproc-per-node N main.py          # one process per GPU
  python main.py --nproc N                     # self-spawned local ranks (no MPI needed)
  python main.py                               # single process

Configuration comes from ``utils.py`` (and ``MPB_*`` environment overrides), as in the
reference; the work itself is ``mpi_pytorch_b200.engine.trainer``.
"""
import argparse
import sys

import utils
from mpi_pytorch_b200.engine.trainer import run_training
from mpi_pytorch_b200.parallel import bootstrap
from mpi_pytorch_b200.utils.config import from_module
from mpi_pytorch_b200.utils.logging import init_logger


def _rank_main(world, settings):
    logger = init_logger('training.log', rank=world.rank, log_dir=settings.LOG_DIR)
    if world.rank == 0:
        logger.info('Logger Initialized')
    return run_training(settings, world)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument('--nproc', type=int, default=0, help='self-spawn this many local ranks')
    args, _ = ap.parse_known_args(argv)
    settings = from_module(utils)
    if args.nproc and args.nproc > 1:
        bootstrap.spawn(_rank_main, args.nproc, (settings,), device=settings.DEVICE)
        return None
    world = bootstrap.init_world(settings.DEVICE)
    try:
        return _rank_main(world, settings)
    finally:
        bootstrap.shutdown()


if __name__ == '__main__':
    main(sys.argv[1:])
