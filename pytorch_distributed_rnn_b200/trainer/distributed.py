"""Shared logic of the synchronous data-parallel trainers (parity: src/motion/trainer/distributed.py).

Only rank 0 keeps the validation / test sets (:20-22); every rank samples its shard with the
epoch-seeded sharded sampler (:35-39); the per-rank batch is ``batch_size // world_size``
(:48-49); log lines carry the rank (:57-58); checkpoints are written by rank 0 only (:60-62).
One process drives one GPU (``cuda:LOCAL_RANK``).
"""
from __future__ import annotations

from ..data.sampler import ShardedEpochSampler
from .base import Trainer
from .formatter import TrainingMessageFormatter


class DistributedTrainer(Trainer):
    def __init__(self, model, training_set, batch_size, learning_rate, rank, world_size,
                 validation_set=None, test_set=None, checkpoint_dir=None, device=None, **kw):
        if rank != 0:
            test_set = None
            validation_set = None
        self.rank = rank
        self.world_size = world_size
        super().__init__(
            model=model, training_set=training_set, validation_set=validation_set, test_set=test_set,
            batch_size=batch_size, learning_rate=learning_rate, checkpoint_dir=checkpoint_dir,
            sampler=ShardedEpochSampler(len(training_set), num_replicas=world_size, rank=rank),
            device=device, **kw)

    def _per_rank_batch(self, batch_size):
        return batch_size // self.world_size

    def _get_formatter(self, epochs):
        return TrainingMessageFormatter(epochs, self.rank)

    def _save_checkpoint(self, epoch, loss, best=False):
        if self.rank == 0:
            super()._save_checkpoint(epoch, loss, best=best)
