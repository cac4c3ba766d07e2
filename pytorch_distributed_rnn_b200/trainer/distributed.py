"""What the synchronous data-parallel trainers have in common.

Capability parity with src/motion/trainer/distributed.py: evaluation data lives on rank 0 only (:20-22),
each rank draws its shard from the epoch-seeded sharded sampler (:35-39), a rank's batch is the global
batch divided by the world size (:48-49), progress lines are prefixed with the rank (:57-58) and only
rank 0 writes checkpoints (:60-62).  Here one process drives one GPU (``cuda:LOCAL_RANK``).
"""
from __future__ import annotations

from ..data.sampler import ShardedEpochSampler
from .base import Trainer
from .formatter import TrainingMessageFormatter


class DistributedTrainer(Trainer):
    def __init__(self, model, training_set, batch_size, learning_rate, rank, world_size, *,
                 validation_set=None, test_set=None, **trainer_kwargs):
        self.rank, self.world_size = int(rank), int(world_size)
        shard = ShardedEpochSampler(len(training_set), num_replicas=self.world_size, rank=self.rank)
        eval_sets = dict(validation_set=validation_set, test_set=test_set) if self.is_root else {}
        # rank 0 alone validates / checkpoints after an epoch; every rank knows whether it will
        self._root_only_work = validation_set is not None or bool(trainer_kwargs.get("save_every"))
        super().__init__(model=model, training_set=training_set, batch_size=batch_size,
                         learning_rate=learning_rate, sampler=shard, **eval_sets, **trainer_kwargs)

    @property
    def is_root(self) -> bool:
        return self.rank == 0

    # -- hooks of the base loop --
    def _per_rank_batch(self, batch_size: int) -> int:
        return batch_size // self.world_size

    def _get_formatter(self, epochs):
        return TrainingMessageFormatter(epochs, self.rank)

    def _epoch_barrier(self) -> None:
        # Host-side barrier: without it the other ranks run ahead into the next epoch's first all-reduce and
        # spin there (against the in-kernel watchdog) for as long as rank 0's evaluation / file write takes.
        if not self._root_only_work or self.world_size == 1:
            return
        import torch
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            if self.device.type == "cuda":
                torch.cuda.synchronize(self.device)
            dist.barrier()

    def _save_checkpoint(self, epoch, loss, best=False):
        if not self.is_root:
            return
        super()._save_checkpoint(epoch, loss, best=best)
