"""``horovod`` sub-command: synchronous data parallelism driven from the optimizer.

Capability parity with src/motion/trainer/horovod.py: ``hvd.init()`` (:17), Adam wrapped in
``hvd.DistributedOptimizer(named_parameters=...)`` (:33-35) and a parameter broadcast from rank 0 before
the first epoch (:40-42).  The model is not wrapped, so checkpoints carry un-prefixed keys.  ``hvd`` is
:mod:`..parallel.horovod` — same call surface, a fixed fusion plan, and on GPUs the very same fused NVLink
all-reduce + optimizer kernel the DDP trainer uses, synchronised at ``step()``.
"""
from __future__ import annotations

from ..parallel import horovod as hvd
from ..parallel import process_group as pg
from ..utils.arena import ParamArena
from .distributed import DistributedTrainer


class HorovodTrainer(DistributedTrainer):
    def __init__(self, model, training_set, batch_size, learning_rate, *, device=None, comm=None, **trainer_kwargs):
        self._comm = hvd.init(comm)
        where = device if device is not None else pg.rank_info().device
        super().__init__(model, training_set, batch_size, learning_rate, hvd.rank(), hvd.size(),
                         device=where, **trainer_kwargs)

    # -- hooks of the base loop --
    def _get_comm(self):
        return self._comm

    def _get_arena(self):
        net = self._core_model()
        make = getattr(self._comm, "make_arena", None)   # symmetric-memory communicators own the parameter arena
        return make(net) if make is not None else ParamArena(net)

    def _get_optimizer(self, model, lr):
        adam = super()._get_optimizer(model, lr)
        return hvd.DistributedOptimizer(adam, named_parameters=model.named_parameters())

    def train(self, epochs):
        hvd.broadcast_parameters(self.model.state_dict(), root_rank=0)   # every rank starts from rank 0's weights
        return super().train(epochs)
