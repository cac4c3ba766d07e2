"""Synthetic motion data of the reference's shape (there is no network, so no UCI-HAR).

Train 6 912 x [128, 9] fp32 with labels ``[N,1]`` int64 in [0,6) — the size every result in
the reference reports ("Training set of size 6912", results_baseline_1.json) — plus
validation (367) and test (2 947) splits of UCI-HAR's sizes.  Sequences are class-dependent
sinusoids + noise so that training actually reduces the loss.  ``write_dataset`` stores them
under the file names ``MotionDataset.load`` (and the reference's dataset.py:71-73) expect.
"""
from __future__ import annotations

from pathlib import Path

import torch

TRAIN_SIZE, VALID_SIZE, TEST_SIZE = 6912, 367, 2947
SEQ_LEN, NUM_FEATURES, NUM_CLASSES = 128, 9, 6


def make_split(n: int, seed: int, seq_len: int = SEQ_LEN, features: int = NUM_FEATURES, classes: int = NUM_CLASSES):
    g = torch.Generator().manual_seed(seed)
    y = torch.randint(0, classes, (n, 1), generator=g)
    t = torch.arange(seq_len, dtype=torch.float32).view(1, seq_len, 1)
    freq = (0.05 + 0.03 * y.float()).view(n, 1, 1)
    phase = torch.rand(n, 1, features, generator=g) * 6.2831853
    amp = 0.5 + 0.1 * torch.arange(features, dtype=torch.float32).view(1, 1, features)
    x = amp * torch.sin(freq * t + phase) + 0.3 * torch.randn(n, seq_len, features, generator=g)
    x = x + 0.2 * (y.float().view(n, 1, 1) - classes / 2)
    return x.contiguous(), y


def make_datasets(seed: int = 0, train: int = TRAIN_SIZE, valid: int = VALID_SIZE, test: int = TEST_SIZE, **kw):
    return make_split(train, seed, **kw), make_split(valid, seed + 1, **kw), make_split(test, seed + 2, **kw)


DONE_MARKER = ".synthetic_complete"


def write_dataset(path, seed: int = 0, **kw) -> Path:
    """Write the six tensors; every file appears atomically (temp name + rename) and a marker file is created
    after the last one, so a concurrent reader never sees a half-written dataset."""
    import os
    path = Path(path)
    path.mkdir(parents=True, exist_ok=True)
    (xtr, ytr), (xv, yv), (xte, yte) = make_datasets(seed, **kw)
    for name, t in (("X_train", xtr), ("y_train", ytr), ("X_validation", xv), ("y_validation", yv),
                    ("X_test", xte), ("y_test", yte)):
        tmp = path / f".{name}.pt.tmp{os.getpid()}"
        torch.save(t, tmp)
        os.replace(tmp, path / f"{name}.pt")
    (path / DONE_MARKER).write_text("ok\n")
    return path


def is_complete(path) -> bool:
    return (Path(path) / DONE_MARKER).exists()
