"""Data path: synthetic shards (determinism, sharding) and the CSV/JPEG file-backed path (C4-C8, C24-C25)."""
import os

import numpy as np
import pytest
import torch

from mpi_pytorch_b200.data.dataset import CSV_COLUMNS, FileShard, GetData, load_index, make_synthetic_index, shard_rows
from mpi_pytorch_b200.data.synthetic import SyntheticShard, synth_images_torch, synth_labels


def test_synthetic_shards_partition_the_dataset():
    total, world = 40, 4
    seen = []
    for r in range(world):
        sh = SyntheticShard(total, 5, 7, (8, 8), r, world, seed=3, shuffle=False)
        assert sh.num_samples == total // world and len(sh) == 2
        seen += sh.indices.tolist()
    assert sorted(seen) == list(range(total))
    a = synth_images_torch(3, torch.tensor([5, 9]), 3, 8, 8)
    b = synth_images_torch(3, torch.tensor([9]), 3, 8, 8)
    assert torch.equal(a[1], b[0]) and a.abs().max() <= 1.7          # sample i depends only on (seed, i)
    assert synth_labels(3, torch.arange(100), 7).max() < 7


def test_shard_rows_array_split_and_equalised():
    parts = [shard_rows(10, r, 3, equalize=False) for r in range(3)]
    assert [len(p) for p in parts] == [4, 3, 3] and np.concatenate(parts).tolist() == list(range(10))
    assert all(len(shard_rows(10, r, 3)) == 3 for r in range(3))      # no uneven-shard hang


def test_file_backed_path(tmp_path):
    tr_csv, te_csv = make_synthetic_index(str(tmp_path), n_train=20, n_test=12, num_classes=5, hw=(30, 22))
    import pandas as pd
    df = pd.read_csv(tr_csv)
    assert list(df.columns[1:]) == CSV_COLUMNS                         # reference schema (data/*.csv:1)
    train, test = load_index(tr_csv, te_csv, debug=False)
    assert len(train) == 20 and len(test) == 12
    dbg_train, dbg_test = load_index(tr_csv, te_csv, debug=True)      # DEBUG samples the *test* csv 80/20
    assert len(dbg_train) + len(dbg_test) == 12
    ds = GetData(str(tmp_path) + "/train/", train["file_name"].values, train["category_id"].values, None)
    img, label = ds[0]
    assert img.size[1] == 30 and img.size[0] in (22, 23) and int(label) == int(train["category_id"].values[0])
    with pytest.raises(ValueError):
        GetData("/somewhere/else/", [], [], None)
    shard = FileShard(str(tmp_path) + "/train/", train, 8, (16, 16), 0, 2, shuffle=False)
    x, y = next(iter(shard))
    assert x.shape == (8, 3, 16, 16) and y.shape == (8,) and shard.num_samples == 10


def test_trainer_on_files(tmp_path):
    from mpi_pytorch_b200.engine.trainer import Trainer
    from mpi_pytorch_b200.utils.config import Settings
    tr_csv, te_csv = make_synthetic_index(str(tmp_path), n_train=16, n_test=8, num_classes=4, hw=(24, 24))
    s = Settings(MODEL_NAME="squeezenet", NUM_CLASSES=4, BATCH_SIZE=8, HEIGHT=32, WIDTH=32, USE_PRETRAINED=False,
                 SYNTHETIC=False, DEBUG=False, TRAIN_CSV=tr_csv, TEST_CSV=te_csv, TRAIN_DIR=str(tmp_path) + "/train/",
                 CHECKPOINT_DIR=str(tmp_path) + "/ck/", MODELS_DIR=str(tmp_path) + "/m/", NUM_EPOCHS=1, DEVICE="cpu")
    hist = Trainer(s).fit()
    assert len(hist) == 1 and np.isfinite(hist[0]["loss"]) and 0.0 <= hist[0]["acc"] <= 1.0
    assert os.path.exists(str(tmp_path) + "/ck/checkpoint_squeezenet.pt")


def test_shipped_index_matches_reference_schema():
    """data/{train,test}_sample.csv ship with the repo (the reference commits its own, data/*.csv:1), so
    SYNTHETIC=False works out of the box; images are rendered on demand from the index rows."""
    import pandas as pd
    from mpi_pytorch_b200.data.dataset import SYNTH_LICENSE, render_index_images
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "data")
    tr, te = pd.read_csv(os.path.join(root, "train_sample.csv")), pd.read_csv(os.path.join(root, "test_sample.csv"))
    assert list(tr.columns[1:]) == CSV_COLUMNS and list(te.columns[1:]) == CSV_COLUMNS
    assert len(tr) == 4000 and len(te) == 1000 and (tr["license"] == SYNTH_LICENSE).all()


def test_mixed_frame_sizes_batch_in_one_launch_per_size(tmp_path):
    from mpi_pytorch_b200.data.dataset import group_by_shape, resize_frames
    frames = [torch.randint(0, 255, (30, 22 + (i % 2), 3), dtype=torch.uint8) for i in range(6)]
    assert group_by_shape(frames) == [[0, 2, 4], [1, 3, 5]]
    from mpi_pytorch_b200 import ops
    x = resize_frames(frames, (16, 16), "cpu", torch.float32, False, "bilinear")
    for i, f in enumerate(frames):
        assert torch.allclose(x[i], ops.resize_normalize(f[None], 16, 16, "bilinear")[0])
