"""UCI-HAR raw-text preprocessor (behavioural parity with src/motion/processor.py).

Reads ``{train,test}/Inertial Signals/<signal>_{train,test}.txt`` for the nine inertial
signals, returns ``FloatTensor [N,128,9]`` features and ``LongTensor [N,1]`` zero-based
labels, splits off a validation fraction with a numpy permutation and truncates the
training set to a multiple of 96 so that 1/2/4/8/12 ranks x 1/2/4 slots divide it evenly
(processor.py:63-66).  Differences, both documented in SURVEY §9: the numpy RNG can be
seeded (Q14), and parsing uses ``numpy.loadtxt`` semantics via ``split()`` instead of the
double-space replace trick (same numbers).
"""
from __future__ import annotations

from pathlib import Path
from typing import Optional

import numpy as np
import torch

TRAIN_MULTIPLE = 96


class MotionDataProcessor:
    TRAIN = "train"
    TEST = "test"
    INPUT_SIGNAL_TYPES = [
        "body_acc_x_", "body_acc_y_", "body_acc_z_",
        "body_gyro_x_", "body_gyro_y_", "body_gyro_z_",
        "total_acc_x_", "total_acc_y_", "total_acc_z_",
    ]

    def __init__(self, seed: Optional[int] = None):
        self.rng = np.random.RandomState(seed) if seed is not None else np.random

    def process_data(self, csv_path, validation_fraction: float = 0.05):
        root = Path(csv_path)
        x_train = self._load_x(root / self.TRAIN / "Inertial Signals", "train")
        x_test = self._load_x(root / self.TEST / "Inertial Signals", "test")
        y_train = self._load_y(root / self.TRAIN / "y_train.txt")
        y_test = self._load_y(root / self.TEST / "y_test.txt")
        (xt, yt), valid = self._train_valid_split(x_train, y_train, validation_fraction)
        n = (len(xt) // TRAIN_MULTIPLE) * TRAIN_MULTIPLE
        return (xt[:n], yt[:n]), valid, (x_test, y_test)

    def _load_x(self, folder: Path, split: str) -> torch.Tensor:
        signals = []
        for sig in self.INPUT_SIGNAL_TYPES:
            with open(folder / f"{sig}{split}.txt", "r") as f:
                signals.append(np.array([[float(v) for v in row.split()] for row in f if row.strip()],
                                        dtype=np.float32))
        # [signals, N, T] -> [N, T, signals]
        return torch.from_numpy(np.ascontiguousarray(np.transpose(np.stack(signals, 0), (1, 2, 0))))

    def _load_y(self, path: Path) -> torch.Tensor:
        with open(path, "r") as f:
            y = np.array([[int(v) for v in row.split()] for row in f if row.strip()], dtype=np.int64)
        return torch.from_numpy(y - 1)

    def _train_valid_split(self, features, labels, validation_fraction):
        n = features.shape[0]
        assert n == labels.shape[0], "Feature and labels must have equal size"
        perm = self.rng.permutation(n)
        k = int(n * validation_fraction)
        vi, ti = torch.from_numpy(perm[:k]), torch.from_numpy(perm[k:])
        return (features[ti], labels[ti]), (features[vi], labels[vi])
