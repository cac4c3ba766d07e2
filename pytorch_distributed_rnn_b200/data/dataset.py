"""In-memory motion dataset + ``.pt`` cache (parity with src/motion/dataset.py).

``MotionDataset.load(base_path, output_path, validation_fraction)`` looks for
``X_{train,validation,test}.pt`` / ``y_*.pt`` under ``base_path`` (dataset.py:38-51,71-73);
if all six exist they are ``torch.load``-ed, otherwise the raw UCI-HAR text files are
preprocessed and the six tensors are written to ``output_path or base_path``.  The cache
files are plain ``torch.save`` tensors, interchangeable with the reference's.

``to(device)`` makes the dataset device-resident so batches are formed by the K11 gather
kernel instead of a host DataLoader.
"""
from __future__ import annotations

import logging
from pathlib import Path
from typing import Optional

import torch
from torch.utils import data

from .processor import MotionDataProcessor


class MotionDataset(data.Dataset):
    LABELS = ["WALKING", "WALKING_UPSTAIRS", "WALKING_DOWNSTAIRS", "SITTING", "STANDING", "LAYING"]

    def __init__(self, features: torch.Tensor, labels: torch.Tensor):
        self.features = features
        self.labels = labels
        self.seq_length = self.features.shape[1]
        self.num_features = self.features.shape[2]

    def __getitem__(self, index):
        return self.features[index], self.labels[index]

    def __len__(self):
        return len(self.features)

    def to(self, device, pin: bool = False) -> "MotionDataset":
        f = self.features.contiguous().float()
        l = self.labels.contiguous().long()
        if pin and torch.cuda.is_available() and not f.is_cuda:
            f, l = f.pin_memory(), l.pin_memory()
        return MotionDataset(f.to(device, non_blocking=True), l.to(device, non_blocking=True))

    @classmethod
    def load(cls, base_path, output_path=None, validation_fraction: float = 0.05, seed: Optional[int] = None):
        base_path = Path(base_path)
        types = ["train", "validation", "test"]
        datasets = []
        for t in types:
            fp, lp = cls.get_data_path(base_path, t)
            if fp.exists() and lp.exists():
                datasets.append(cls(torch.load(fp), torch.load(lp)))
        if len(datasets) == 3:
            logging.info("Preprocessed data found. Skip preprocessing.")
            return datasets
        output_path = base_path if output_path is None else Path(output_path)
        logging.info("No processed data found. Preprocess raw data...")
        (xtr, ytr), (xv, yv), (xte, yte) = MotionDataProcessor(seed).process_data(base_path, validation_fraction)
        output_path.mkdir(parents=True, exist_ok=True)
        for name, t in (("X_train", xtr), ("y_train", ytr), ("X_validation", xv), ("y_validation", yv),
                        ("X_test", xte), ("y_test", yte)):
            torch.save(t, output_path / f"{name}.pt")
        return cls(xtr, ytr), cls(xv, yv), cls(xte, yte)

    @classmethod
    def get_data_path(cls, base_path, data_type):
        base_path = Path(base_path)
        return base_path / f"X_{data_type}.pt", base_path / f"y_{data_type}.pt"
