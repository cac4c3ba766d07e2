from .dataset import MotionDataset
from .processor import MotionDataProcessor
from .sampler import ShardedEpochSampler, epoch_indices
from . import synthetic

__all__ = ["MotionDataset", "MotionDataProcessor", "ShardedEpochSampler", "epoch_indices", "synthetic"]
