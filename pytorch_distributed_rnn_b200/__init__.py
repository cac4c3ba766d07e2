"""B200-native distributed RNN training (capabilities of jkhlr/pytorch-distributed-rnn).

Layers: ``csrc/`` + ``include/b200rnn.h`` (hand-written sm_100a kernels, C ABI, C++ runtime
bootstrap) -> :mod:`._native` (ctypes) -> :mod:`.ops` (autograd glue, fused optimizers, the
3-launch training engine) -> :mod:`.models` / :mod:`.data` -> :mod:`.parallel` (process
group, DDP / Horovod-style / parameter-server strategies, NVLink symmetric-memory
communicator) -> :mod:`.trainer` -> :mod:`.cli`.
"""
from . import _native
from .data import MotionDataset, ShardedEpochSampler, epoch_indices
from .models import FusedGRU, FusedLSTM, MotionModel, ToyDataset, ToyModel
from .ops.optim import FusedAdam, FusedSGD
from .utils.arena import ParamArena, params_checksum

__version__ = "0.1.0"

__all__ = ["MotionModel", "FusedLSTM", "FusedGRU", "ToyModel", "ToyDataset", "MotionDataset",
           "ShardedEpochSampler", "epoch_indices", "FusedAdam", "FusedSGD", "ParamArena",
           "params_checksum", "native_available"]


def native_available() -> bool:
    return _native.available()
