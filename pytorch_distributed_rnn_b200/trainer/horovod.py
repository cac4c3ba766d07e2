"""``horovod`` sub-command: optimizer-wrapped synchronous DP (parity: src/motion/trainer/horovod.py).

``hvd.init()`` (:17), Adam wrapped in ``hvd.DistributedOptimizer(named_parameters=...)``
(:33-35), and ``hvd.broadcast_parameters(model.state_dict(), root_rank=0)`` at the top of
``train()`` (:40-42).  The model is not wrapped, so checkpoints have un-prefixed keys.
``hvd`` is :mod:`..parallel.horovod` — same API, fixed fusion plan, and on GPUs the very same
fused NVLink all-reduce + optimizer kernel as the DDP trainer, synchronised at ``step()``.
"""
from __future__ import annotations

from ..parallel import horovod as hvd
from ..utils.arena import ParamArena
from .distributed import DistributedTrainer


class HorovodTrainer(DistributedTrainer):
    def __init__(self, model, training_set, batch_size, learning_rate, validation_set=None, test_set=None,
                 checkpoint_dir=None, device=None, comm=None, **kw):
        self._comm = hvd.init(comm)
        from ..parallel import process_group as pg
        info = pg.rank_info()
        super().__init__(rank=hvd.rank(), world_size=hvd.size(), model=model, training_set=training_set,
                         validation_set=validation_set, test_set=test_set, batch_size=batch_size,
                         learning_rate=learning_rate, checkpoint_dir=checkpoint_dir,
                         device=device if device is not None else info.device, **kw)

    def _get_arena(self):
        m = self._core_model()
        return self._comm.make_arena(m) if hasattr(self._comm, "make_arena") else ParamArena(m)

    def _get_optimizer(self, model, lr):
        optimizer = super()._get_optimizer(model, lr)
        return hvd.DistributedOptimizer(optimizer, named_parameters=model.named_parameters())

    def _get_comm(self):
        return self._comm

    def train(self, epochs):
        hvd.broadcast_parameters(self.model.state_dict(), root_rank=0)
        return super().train(epochs)
