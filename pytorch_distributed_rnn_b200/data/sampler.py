"""Epoch-seeded sharded sampling, index-compatible with torch's ``DistributedSampler``.

The reference uses ``DistributedSampler`` for every trainer — even ``local`` runs use
``num_replicas=1, rank=0`` (src/motion/trainer/base.py:22-26, distributed.py:35-39) — and
re-seeds it per epoch (base.py:74-75).  That is what makes the union of the W per-rank
batches at step k equal the local batch at step k.  Here the permutation is
computed once per epoch as one index vector (27 KB for 6 912 samples) that is uploaded to
the GPU and consumed by the K11 gather kernel; no per-sample Python ``__getitem__``.
"""
from __future__ import annotations

import math

import torch


def epoch_indices(n: int, num_replicas: int = 1, rank: int = 0, epoch: int = 0, seed: int = 0,
                  shuffle: bool = True, drop_last: bool = False) -> torch.Tensor:
    """The exact index sequence ``DistributedSampler(ds, num_replicas, rank, shuffle, seed, drop_last)``
    yields after ``set_epoch(epoch)`` (int64 CPU tensor)."""
    if shuffle:
        g = torch.Generator()
        g.manual_seed(seed + epoch)
        idx = torch.randperm(n, generator=g)
    else:
        idx = torch.arange(n)
    if drop_last and n % num_replicas != 0:
        num_samples = math.ceil((n - num_replicas) / num_replicas)
    else:
        num_samples = math.ceil(n / num_replicas)
    total = num_samples * num_replicas
    if not drop_last:
        pad = total - idx.numel()
        if pad > 0:
            reps = math.ceil(pad / idx.numel())
            idx = torch.cat([idx, idx.repeat(reps)[:pad]])
    else:
        idx = idx[:total]
    return idx[rank:total:num_replicas].contiguous()


class ShardedEpochSampler:
    """Stateful wrapper with the ``set_epoch`` / ``__iter__`` / ``__len__`` surface of DistributedSampler."""

    def __init__(self, n: int, num_replicas: int = 1, rank: int = 0, seed: int = 0, shuffle: bool = True):
        self.n, self.num_replicas, self.rank, self.seed, self.shuffle = n, num_replicas, rank, seed, shuffle
        self.epoch = 0

    def set_epoch(self, epoch: int) -> None:
        self.epoch = epoch

    def indices(self) -> torch.Tensor:
        return epoch_indices(self.n, self.num_replicas, self.rank, self.epoch, self.seed, self.shuffle)

    def __iter__(self):
        return iter(self.indices().tolist())

    def __len__(self) -> int:
        return math.ceil(self.n / self.num_replicas)
