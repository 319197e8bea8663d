"""Real-data path: CSV index, DEBUG sampling, sharding and the file-backed Dataset.

Covers the reference's C4-C8 / C24-C25: ``GetData`` (data_loader.py:6-39),
the rank-0 index load with DEBUG sampling (main.py:73-82), sharding (main.py:84,91) and
the dataset sampling script (create_dataset.py).  The heavy lifting the reference does
with PIL on the CPU per sample (resize + normalise) is done on the GPU by the fused
``ops.resize_normalize`` kernel after the uint8 image has been uploaded.
"""
from __future__ import annotations

import json
import os
import shutil
from typing import Any, List, Optional, Sequence, Tuple

import numpy as np
import torch
from torch.utils.data import Dataset

CSV_COLUMNS = ["file_name", "height", "id", "license", "width", "category_id", "institution_id"]


class GetData(Dataset):
    """``GetData(Dir, FNames, Labels, Transform)`` as in data_loader.py:10.

    Returns ``(image, label)`` for a training directory and ``(image, file_name)`` for a
    test directory.  The reference decides by substring match on ``Dir`` and silently
    returns ``None`` when neither matches; here ``mode`` can be given
    explicitly and an undecidable directory raises."""

    def __init__(self, Dir: str, FNames: Sequence[str], Labels: Sequence[int], Transform: Any = None,
                 mode: Optional[str] = None):
        self.dir, self.fnames, self.labels, self.transform = Dir, FNames, Labels, Transform
        if mode is None:
            mode = "train" if "train" in Dir else ("test" if "test" in Dir else None)
        if mode not in ("train", "test"):
            raise ValueError("cannot infer train/test from directory {!r}; pass mode=".format(Dir))
        self.mode = mode

    def __len__(self) -> int:
        return len(self.fnames)

    def load_uint8(self, index: int) -> torch.Tensor:
        from PIL import Image
        with Image.open(os.path.join(self.dir, self.fnames[index])) as im:
            return torch.from_numpy(np.asarray(im.convert("RGB")).copy())      # HWC uint8

    def __getitem__(self, index: int):
        from PIL import Image
        x = Image.open(os.path.join(self.dir, self.fnames[index]))
        if self.transform is not None:
            x = self.transform(x)
        return (x, self.labels[index]) if self.mode == "train" else (x, self.fnames[index])


def load_index(train_csv: str, test_csv: str, debug: bool, seed: int = 0):
    """Rank-0 index load of main.py:73-82.  DEBUG: 1000 rows of the *test* csv
    (``random_state=0``) split 80/20; the split is seeded here (the reference's is not)."""
    import pandas as pd
    from sklearn.model_selection import train_test_split
    if debug:
        df = pd.read_csv(test_csv)
        sample = df.sample(min(1000, len(df)), random_state=0).reset_index(drop=True)
        train, test = train_test_split(sample, test_size=0.2, random_state=seed)
        return train.reset_index(drop=True), test.reset_index(drop=True)
    return pd.read_csv(train_csv), pd.read_csv(test_csv)


def shard_rows(n_rows: int, rank: int, world: int, equalize: bool = True) -> np.ndarray:
    """Row ids owned by ``rank``.  ``np.array_split`` semantics (contiguous, sizes differ by at
    most one; main.py:84) unless ``equalize`` truncates every shard to the common minimum so
    all ranks run the same number of steps."""
    parts = np.array_split(np.arange(n_rows), world)
    mine = parts[rank]
    if equalize:
        mine = mine[: min(len(p) for p in parts)]
    return mine


def read_json(path: str) -> Any:
    """create_dataset.py:17-24 (Herbarium metadata is ISO-8859-1)."""
    with open(path, "r", encoding="ISO-8859-1") as fh:
        return json.load(fh)


def create_dataframe(ann: Any):
    """images JOIN annotations on id (create_dataset.py:34-39)."""
    import pandas as pd
    img = pd.DataFrame(ann["images"])
    an = pd.DataFrame(ann["annotations"]).drop(columns="image_id")
    return img.merge(an, on="id")


def create_dataset(train_dir: str, train_file: str, n_images: int, out_dir: str = "./data", copy_images: bool = True,
                   seed: int = 0) -> Tuple[str, str]:
    """Sample N images, 80/20 split, write the two CSVs, optionally copy image files
    (create_dataset.py:42-66)."""
    from sklearn.model_selection import train_test_split
    df = create_dataframe(read_json(os.path.join(train_dir, train_file)))
    sample = df.sample(min(n_images, len(df)), random_state=0).reset_index(drop=True)
    train, test = train_test_split(sample, test_size=0.2, random_state=seed)
    os.makedirs(out_dir, exist_ok=True)
    p_tr, p_te = os.path.join(out_dir, "train_sample.csv"), os.path.join(out_dir, "test_sample.csv")
    train.to_csv(p_tr)
    test.to_csv(p_te)
    if copy_images:
        for mode, part in (("test", test), ("train", train)):
            for src in part.file_name:
                dst = os.path.join(out_dir, "img", mode, src)
                os.makedirs(os.path.dirname(dst), exist_ok=True)
                shutil.copy2(os.path.join(train_dir, src), dst)
    return p_tr, p_te
