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


def group_by_shape(frames: Sequence[torch.Tensor]):
    """Positions of ``frames`` grouped by frame shape, in first-seen order."""
    groups: dict = {}
    for j, f in enumerate(frames):
        groups.setdefault(tuple(f.shape), []).append(j)
    return list(groups.values())


def resize_frames(frames: Sequence[torch.Tensor], hw: Tuple[int, int], device, dtype: torch.dtype,
                  channels_last: bool, mode: str) -> torch.Tensor:
    """Resize + normalise a list of HWC uint8 frames: ONE pinned upload and ONE kernel launch per distinct frame
    size (every image is resized on its own, as ``pil_image.resize`` / ``transforms.Resize`` do in the reference)."""
    from .. import ops
    device = torch.device(device)
    out: Optional[torch.Tensor] = None
    for pos in group_by_shape(frames):
        stack = torch.stack([frames[j] for j in pos])
        if device.type == "cuda":
            stack = stack.pin_memory().to(device, non_blocking=True)
        y = ops.resize_normalize(stack, hw[0], hw[1], mode, dtype, channels_last)
        if len(pos) == len(frames):
            return y
        if out is None:
            out = torch.empty((len(frames),) + tuple(y.shape[1:]), dtype=y.dtype, device=y.device)
            if channels_last:
                out = out.contiguous(memory_format=torch.channels_last)
        out[torch.as_tensor(pos, device=y.device)] = y
    return out


class FileShard:
    """This rank's slice of a CSV-indexed JPEG dataset, batched for the GPU.

    The reference decodes, resizes and normalises every sample on the CPU inside a 0-worker
    DataLoader (main.py:62-65,102; data_loader.py:29-37).  Here the CPU only decodes to uint8; each
    frame is uploaded as-is and resized + normalised on the device by the fused ``ops.resize_normalize``
    kernel (train-time semantics: antialiased bilinear on the float image, then ImageNet mean/std)."""

    def __init__(self, directory: str, frame, batch_size: int, hw: Tuple[int, int], rank: int = 0, world: int = 1,
                 device="cpu", dtype: torch.dtype = torch.float32, channels_last: bool = False, shuffle: bool = True,
                 seed: int = 0, mode: str = "train"):
        rows = shard_rows(len(frame), rank, world, equalize=True)
        part = frame.iloc[rows]
        self.ds = GetData(directory, part["file_name"].values, part["category_id"].values, None, mode=mode)
        self.batch_size, self.hw, self.device = int(batch_size), hw, torch.device(device)
        self.dtype, self.channels_last, self.shuffle, self.seed, self.rank = dtype, channels_last, shuffle, seed, rank
        self.total = len(frame)
        self.epoch = 0

    def __len__(self) -> int:
        return (len(self.ds) + self.batch_size - 1) // self.batch_size

    @property
    def num_samples(self) -> int:
        return len(self.ds)

    def set_epoch(self, epoch: int) -> None:
        self.epoch = int(epoch)

    def __iter__(self):
        n = len(self.ds)
        order = torch.arange(n)
        if self.shuffle:
            g = torch.Generator().manual_seed(self.seed * 1000003 + self.epoch * 7919 + self.rank)
            order = torch.randperm(n, generator=g)
        for b in range(len(self)):
            idx = order[b * self.batch_size:(b + 1) * self.batch_size].tolist()
            frames = [self.ds.load_uint8(i) for i in idx]
            labels = torch.as_tensor([int(self.ds.labels[i]) for i in idx], dtype=torch.int64)
            x = resize_frames(frames, self.hw, self.device, self.dtype, self.channels_last, "bilinear")
            if self.channels_last:
                x = x.contiguous(memory_format=torch.channels_last)
            yield x, labels.to(self.device)


SYNTH_LICENSE = 999       # `license` value that marks rows of a generated (non-Herbarium) index


def _render_row(directory: str, rel: str, h: int, w: int, cat: int, sid: int) -> None:
    """Deterministic stand-in JPEG for one index row: noise plus a class-dependent colour cast, so a model can
    actually learn the labels."""
    from PIL import Image
    full = os.path.join(directory, rel)
    if os.path.exists(full):
        return
    os.makedirs(os.path.dirname(full), exist_ok=True)
    rng = np.random.RandomState(sid % (2 ** 31))
    arr = (rng.rand(h, w, 3) * 96).astype(np.int32)
    arr[:, :, cat % 3] += 40 + 8 * (cat % 16)
    arr[: h // 2, :, (cat // 3) % 3] += 6 * (cat % 24)
    Image.fromarray(np.clip(arr, 0, 255).astype(np.uint8)).save(full, quality=95)


def render_index_images(frame, directory: str) -> int:
    """Write the JPEG of every generated row of ``frame`` that is missing under ``directory``; rows of a real
    (Herbarium) index are left alone.  Returns the number of generated rows."""
    gen = frame[frame["license"] == SYNTH_LICENSE]
    for rel, h, w, cat, sid in zip(gen["file_name"], gen["height"], gen["width"], gen["category_id"], gen["id"]):
        _render_row(directory, rel, int(h), int(w), int(cat), int(sid))
    return len(gen)


def make_synthetic_index(out_dir: str, n_train: int = 64, n_test: int = 16, num_classes: int = 10, hw=(40, 28),
                         seed: int = 0, render: bool = True, csv_dir: Optional[str] = None):
    """Write a small Herbarium-shaped dataset: the two CSV indices with the reference's columns (data/*.csv:1) and,
    with ``render``, the JPEG files under ``out_dir/{train,test}/`` -- so the file-backed path runs without the 6 GB
    original.  Frame widths vary by a few pixels like the real set (680/681 there)."""
    import pandas as pd
    rng = np.random.RandomState(seed)
    paths = {}
    csv_dir = csv_dir or out_dir
    os.makedirs(csv_dir, exist_ok=True)
    nid = 100000
    for split, n in (("train", n_train), ("test", n_test)):
        rows = []
        for i in range(n):
            cat = int(rng.randint(0, num_classes))
            w = hw[1] + int(rng.randint(0, 2))
            rel = "images/{:03d}/{:02d}/{}.jpg".format(cat // 100, cat % 100, nid)
            rows.append({"file_name": rel, "height": hw[0], "id": nid, "license": SYNTH_LICENSE, "width": w,
                         "category_id": cat, "institution_id": 0})
            nid += 1
        df = pd.DataFrame(rows, columns=CSV_COLUMNS)
        paths[split] = os.path.join(csv_dir, "{}_sample.csv".format(split))
        df.to_csv(paths[split])
        if render:
            # both splits are read from the *training* image directory, as in the reference
            # (main.py:99-109 and evaluation_pipeline.py:59 use TRAIN_DIR for test_sample.csv rows too)
            render_index_images(df, os.path.join(out_dir, "train"))
            if split == "test":
                render_index_images(df, os.path.join(out_dir, "test"))
    return paths["train"], paths["test"]
