"""Toy MLP + dataset of the DDP plumbing example (src/example/example_ddp.py:11-32).

``ToyModel`` = Linear(10,10) -> ReLU -> Linear(10,5) (165 parameters = 660 B);
``ToyDataset(24)`` = randn features [24,10] and labels [24,5].
"""
from __future__ import annotations

import torch
from torch import nn
from torch.utils.data import Dataset


class ToyModel(nn.Module):
    def __init__(self):
        super().__init__()
        self.net1 = nn.Linear(10, 10)
        self.relu = nn.ReLU()
        self.net2 = nn.Linear(10, 5)

    def forward(self, x):
        return self.net2(self.relu(self.net1(x)))


class ToyDataset(Dataset):
    def __init__(self, size: int):
        self.size = size
        self.features = torch.randn(size, 10)
        self.labels = torch.randn(size, 5)

    def __getitem__(self, index):
        return self.features[index], self.labels[index]

    def __len__(self):
        return self.size
