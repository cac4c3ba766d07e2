from .motion import MotionModel
from .rnn import FusedGRU, FusedLSTM, rnn_composite
from .toy import ToyDataset, ToyModel

__all__ = ["MotionModel", "FusedLSTM", "FusedGRU", "rnn_composite", "ToyModel", "ToyDataset"]
