"""``distributed`` sub-command: DistributedDataParallel trainer (parity: src/motion/trainer/ddp.py).

The reference does ``init_process_group('mpi')`` + ``DistributedDataParallel(model)`` and lets
c10d's Reducer all-reduce one 56.6 KB bucket per step over MPI.  Here the
process group comes from :mod:`..parallel.process_group`, the wrapper from
:mod:`..parallel.ddp` (constructor broadcast, ``module.``-prefixed state_dict), and on GPUs
the gradient all-reduce + averaging + Adam step is a single hand-written kernel over NVLink
peer memory launched right behind the backward kernel (:mod:`..parallel.symm`).
"""
from __future__ import annotations

from ..parallel import process_group as pg
from ..parallel.comm import make_comm
from ..parallel.ddp import DistributedDataParallel
from ..utils.arena import ParamArena
from .distributed import DistributedTrainer


class DDPTrainer(DistributedTrainer):
    def __init__(self, model, training_set, batch_size, learning_rate, validation_set=None, test_set=None,
                 checkpoint_dir=None, device=None, comm=None, **kw):
        info = pg.init_distributed()
        self._info = info
        self._comm = comm if comm is not None else make_comm(info)
        super().__init__(rank=info.rank, world_size=info.world_size, model=model, training_set=training_set,
                         validation_set=validation_set, test_set=test_set, batch_size=batch_size,
                         learning_rate=learning_rate, checkpoint_dir=checkpoint_dir,
                         device=device if device is not None else info.device, **kw)

    def _place_model(self, model):
        model = model.to(self.device)
        arena = self._comm.make_arena(model) if hasattr(self._comm, "make_arena") else ParamArena(model)
        self._ddp = DistributedDataParallel(model, comm=self._comm, arena=arena)
        return self._ddp

    def _core_model(self):
        return self._ddp.module

    def _get_arena(self):
        return self._ddp.arena

    def _get_comm(self):
        return self._comm
