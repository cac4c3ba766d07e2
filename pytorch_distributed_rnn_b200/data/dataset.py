"""In-memory motion dataset and its on-disk tensor cache.

Capability parity with src/motion/dataset.py: ``MotionDataset.load(base_path, output_path, validation_fraction)``
returns the (train, validation, test) datasets, from the six cached tensors ``X_<split>.pt`` / ``y_<split>.pt``
under ``base_path`` when all of them exist (dataset.py:38-51,71-73), otherwise by preprocessing the raw UCI-HAR
text files and writing the cache to ``output_path or base_path``.  The cache files are plain ``torch.save``
tensors, interchangeable with the reference's.

Beyond the reference: ``to(device)`` makes a dataset device-resident, so training batches are formed by the K11
gather kernel from an index vector instead of by a host DataLoader; ``seed`` makes the validation split
reproducible (quirk Q14).
"""
from __future__ import annotations

import logging
from pathlib import Path
from typing import Dict, Optional, Tuple

import torch
from torch.utils import data

from .processor import MotionDataProcessor

SPLITS = ("train", "validation", "test")


class _TensorCache:
    """The six ``.pt`` files of one dataset directory."""

    def __init__(self, directory):
        self.directory = Path(directory)

    def paths(self, split: str) -> Tuple[Path, Path]:
        return self.directory / f"X_{split}.pt", self.directory / f"y_{split}.pt"

    def complete(self) -> bool:
        return all(p.exists() for split in SPLITS for p in self.paths(split))

    def read(self) -> Dict[str, Tuple[torch.Tensor, torch.Tensor]]:
        return {split: tuple(torch.load(p) for p in self.paths(split)) for split in SPLITS}

    def write(self, tensors: Dict[str, Tuple[torch.Tensor, torch.Tensor]]) -> None:
        self.directory.mkdir(parents=True, exist_ok=True)
        for split, pair in tensors.items():
            for path, tensor in zip(self.paths(split), pair):
                torch.save(tensor, path)


class MotionDataset(data.Dataset):
    LABELS = ["WALKING", "WALKING_UPSTAIRS", "WALKING_DOWNSTAIRS", "SITTING", "STANDING", "LAYING"]

    def __init__(self, features: torch.Tensor, labels: torch.Tensor):
        assert len(features) == len(labels), "features and labels differ in length"
        self.features, self.labels = features, labels
        _, self.seq_length, self.num_features = features.shape

    def __len__(self) -> int:
        return self.features.shape[0]

    def __getitem__(self, index):
        return self.features[index], self.labels[index]

    def to(self, device, pin: bool = False) -> "MotionDataset":
        """A copy resident on ``device`` (fp32 features, int64 labels, contiguous: what the gather kernel reads)."""
        feats, labels = self.features.contiguous().float(), self.labels.contiguous().long()
        if pin and torch.cuda.is_available() and not feats.is_cuda:
            feats, labels = feats.pin_memory(), labels.pin_memory()
        return MotionDataset(feats.to(device, non_blocking=True), labels.to(device, non_blocking=True))

    @classmethod
    def get_data_path(cls, base_path, data_type: str) -> Tuple[Path, Path]:
        return _TensorCache(base_path).paths(data_type)

    @classmethod
    def load(cls, base_path, output_path=None, validation_fraction: float = 0.05, seed: Optional[int] = None):
        cache = _TensorCache(base_path)
        if cache.complete():
            logging.info("Preprocessed data found. Skip preprocessing.")
            tensors = cache.read()
        else:
            logging.info("No processed data found. Preprocess raw data...")
            parts = MotionDataProcessor(seed).process_data(Path(base_path), validation_fraction)
            tensors = dict(zip(SPLITS, parts))
            _TensorCache(base_path if output_path is None else output_path).write(tensors)
        return tuple(cls(*tensors[split]) for split in SPLITS)
