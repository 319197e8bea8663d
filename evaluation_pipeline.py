"""Pipelined evaluation entry point (reference surface: /root/reference/evaluation_pipeline.py).

Same launch forms as ``main.py``.  Rank roles for N >= 4: 0 reads, 1 resizes, 2
normalises + routes, 3..N-1 predict with the checkpoint written by ``main.py``; fewer
ranks fold the stages (``mpi_pytorch_b200.parallel.pipeline``).
"""
import argparse
import os
import sys

import torch

import utils
from mpi_pytorch_b200.data.synthetic import synth_labels
from mpi_pytorch_b200.engine.checkpoint import checkpoint_path
from mpi_pytorch_b200.models import initialize_model
from mpi_pytorch_b200.parallel import bootstrap
from mpi_pytorch_b200.parallel.pipeline import EvalPipeline
from mpi_pytorch_b200.utils.config import from_module
from mpi_pytorch_b200.utils.logging import init_logger

RAW_HW = (250, 170)      # synthetic stand-in for the ~1000x680 Herbarium JPEGs, same aspect ratio


def build_sources(settings, world, n):
    """Synthetic raw uint8 frames + labels, or the CSV/JPEG path when SYNTHETIC is off."""
    if settings.SYNTHETIC:
        labels = synth_labels(settings.SEED, torch.arange(n), settings.NUM_CLASSES)

        def source(i0, i1):
            g = torch.Generator().manual_seed(settings.SEED * 7919 + i0)
            return torch.randint(0, 256, (i1 - i0, RAW_HW[0], RAW_HW[1], 3), generator=g, dtype=torch.uint8)
        return source, labels, RAW_HW, None
    import pandas as pd
    from mpi_pytorch_b200.data.dataset import GetData, render_index_images
    df = pd.read_csv(settings.TEST_CSV).iloc[:n]
    if world.rank == 0:
        render_index_images(df, settings.TEST_DIR)       # stand-in JPEGs for generated rows; no-op for Herbarium
    ds = GetData(settings.TEST_DIR, df['file_name'].values, df['category_id'].values, None, mode='test')
    hmax, wmax = int(df['height'].max()), int(df['width'].max())
    frame_hw = torch.as_tensor(df[['height', 'width']].values.copy())

    def source(i0, i1):
        # a list of frames of their own sizes: stage 1 resizes each image over its true extent, as
        # pil_image.resize does in the reference (evaluation_pipeline.py:89), not over a padded canvas
        return [ds.load_uint8(i) for i in range(i0, i1)]
    return source, torch.as_tensor(df['category_id'].values.copy()), (hmax, wmax), frame_hw


def pipeline(world=None, settings=None, n_samples=None):
    settings = settings or from_module(utils)
    world = world or bootstrap.init_world(settings.DEVICE)
    logger = init_logger('evaluation.log', rank=world.rank, log_dir=settings.LOG_DIR)
    if world.rank == 0:
        logger.info('Logger Initialized')
    n = n_samples or (1000 if settings.DEBUG else 10000)
    source, labels, raw_hw, frame_hw = build_sources(settings, world, n)
    n = min(n, int(labels.numel()))

    def model_fn():
        model, _ = initialize_model(settings.MODEL_NAME, settings.NUM_CLASSES, use_pretrained=False,
                                    feature_extract=False)
        path = checkpoint_path(settings.CHECKPOINT_DIR, settings.MODEL_NAME)
        if os.path.exists(path):
            model.load_state_dict(torch.load(path, map_location='cpu', weights_only=True)['state_dict'])
        else:
            logger.warning('no checkpoint at {}: predicting with random weights'.format(path))
        if world.device.type == 'cuda':
            from mpi_pytorch_b200 import ops
            model = model.to(world.device).to(memory_format=torch.channels_last)
            # parameters and BatchNorm statistics stay fp32; GEMM-shaped weights get a bf16 compute copy
            model = ops.prepare_inference(model, torch.bfloat16 if settings.DTYPE == 'bf16' else torch.float32)
        return model
    dtype = torch.bfloat16 if (world.device.type == 'cuda' and settings.DTYPE == 'bf16') else torch.float32
    coll = None
    if world.device.type == 'cuda' and world.size > 1:
        from mpi_pytorch_b200.parallel.collectives import make_collectives
        coll = make_collectives(world, settings.ALLREDUCE_ALGO)     # final count: mpb_count_allreduce over NVLink
    pipe = EvalPipeline(world, n, raw_hw, (settings.HEIGHT, settings.WIDTH), source, labels, model_fn,
                        micro_batch=settings.PIPE_MICROBATCH, seed=settings.SEED, dtype=dtype, collectives=coll,
                        frame_hw=frame_hw)
    return pipe.run()


def _rank_main(world, settings):
    return pipeline(world, settings)


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__, formatter_class=argparse.RawDescriptionHelpFormatter)
    ap.add_argument('--nproc', type=int, default=0)
    args, _ = ap.parse_known_args(argv)
    settings = from_module(utils)
    if args.nproc and args.nproc > 1:
        bootstrap.spawn(_rank_main, args.nproc, (settings,), device=settings.DEVICE)
        return None
    try:
        return pipeline(None, settings)
    finally:
        bootstrap.shutdown()


if __name__ == '__main__':
    main(sys.argv[1:])
