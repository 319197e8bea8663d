"""Run constants -- edit and re-run, exactly like the reference's ``utils.py``.

Same names and defaults as /root/reference/utils.py:4-45 (nothing removed), plus the new
B200 knobs documented in ``mpi_pytorch_b200/utils/config.py``.  Paths are relative to the
working directory instead of the reference's hard-coded ``./project/project_git/...``
(SURVEY D11).  Any value can be overridden without editing this file through an
``MPB_<NAME>`` environment variable.
"""
from mpi_pytorch_b200.utils.config import Settings as _Settings, from_module as _from_module

_S = _from_module(mod=False)          # defaults + MPB_* environment overrides

MODEL_NAME = _S.MODEL_NAME            # resnet18 | resnet34 | alexnet | vgg | vgg16 | squeezenet | densenet | inception
FROM_CHECKPOINT = _S.FROM_CHECKPOINT  # resume from ./checkpoints/checkpoint_<model>.pt
VALIDATE = _S.VALIDATE                # rank 0 measures accuracy after every epoch
CHECKPOINT_NAME = 'checkpoint_{}.pt'.format(MODEL_NAME)

DEBUG = _S.DEBUG                      # small 1000-row sample (800 train rows)
N_IMAGES = _S.N_IMAGES

TRAIN_DIR = _S.TRAIN_DIR
TRAIN_FILE = _S.TRAIN_FILE
TEST_DIR = _S.TEST_DIR
CHECKPOINT_DIR = _S.CHECKPOINT_DIR
MODELS_DIR = _S.MODELS_DIR

WIDTH = _S.WIDTH
HEIGHT = _S.HEIGHT

NUM_CLASSES = _S.NUM_CLASSES
BATCH_SIZE = _S.BATCH_SIZE
LR = _S.LR
NUM_EPOCHS = _S.NUM_EPOCHS
FEATURE_EXTRACT = _S.FEATURE_EXTRACT
USE_PRETRAINED = _S.USE_PRETRAINED

# ---- B200 runtime knobs ------------------------------------------------------------
DEVICE = _S.DEVICE
DTYPE = _S.DTYPE
SYNTHETIC = _S.SYNTHETIC
SYNTHETIC_SAMPLES = _S.SYNTHETIC_SAMPLES
USE_INPUT_SIZE = _S.USE_INPUT_SIZE
OPTIMIZER = _S.OPTIMIZER
MOMENTUM = _S.MOMENTUM
BUCKET_MB = _S.BUCKET_MB
ALLREDUCE_ALGO = _S.ALLREDUCE_ALGO
GRAD_DTYPE = _S.GRAD_DTYPE
NATIVE_OPS = _S.NATIVE_OPS
USE_CUDA_GRAPH = _S.USE_CUDA_GRAPH
SEED = _S.SEED
TRAIN_CSV = _S.TRAIN_CSV
TEST_CSV = _S.TEST_CSV
LOG_DIR = _S.LOG_DIR
PIPE_MICROBATCH = _S.PIPE_MICROBATCH
