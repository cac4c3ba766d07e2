"""Motion sequence classifier: stacked LSTM (or GRU) -> last timestep -> Linear.

Same module tree and state_dict keys as the reference's ``MotionModel``
(src/motion/model.py:4-17): ``lstm.weight_ih_l0`` ... ``lstm.bias_hh_l{L-1}``,
``fc.weight [C,H]``, ``fc.bias [C]``.  The output is raw logits (no softmax), the
initial (h, c) is zero on every call, and ``--dropout`` is accepted by the CLI but not
used (reference quirk Q1).  With the same ``torch.manual_seed`` the initial weights are
bit-identical to the reference's, because parameters are created and initialised in the
same order from the same RNG stream.
"""
from __future__ import annotations

import torch
from torch import nn

from .rnn import FusedGRU, FusedLSTM


class MotionModel(nn.Module):
    def __init__(self, input_dim: int, hidden_dim: int, layer_dim: int, output_dim: int, cell: str = "lstm"):
        super().__init__()
        self.input_dim = input_dim
        self.hidden_dim = hidden_dim
        self.layer_dim = layer_dim
        self.output_dim = output_dim
        self.cell = cell
        rnn_cls = {"lstm": FusedLSTM, "gru": FusedGRU}[cell]
        # attribute is called `lstm` for both cells so checkpoints keep the reference's key names
        self.lstm = rnn_cls(input_dim, hidden_dim, layer_dim, batch_first=True)
        self.fc = nn.Linear(hidden_dim, output_dim)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        out, _ = self.lstm(x)
        return self.fc(out[:, -1, :])
