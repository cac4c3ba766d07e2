"""LSTM / GRU modules with torch.nn.LSTM-compatible parameters and state_dict.

``FusedLSTM(input, hidden, layers, batch_first=True)`` is a drop-in for the
``nn.LSTM`` the reference builds (src/motion/model.py:9): same parameter names
(``weight_ih_l{k}`` ``[4H,in]``, ``weight_hh_l{k}`` ``[4H,H]``, ``bias_ih_l{k}``,
``bias_hh_l{k}``; gate order i,f,g,o), same initialisation stream (so the same
``torch.manual_seed`` gives bit-identical initial weights to stock torch), same
call signature and return values.  What differs is what runs underneath:

* CUDA tensors -> the hand-written sm_100a kernels in ``csrc/rnn``: the persistent fused family
  (:mod:`..ops.rnn_function`; H = 32 LSTM, one launch for the whole sequence and all layers) or the
  large-hidden path on the tcgen05 GEMM (:mod:`..ops.rnn_large`; LSTM and GRU, H % 8 == 0); only
  shapes outside both (H not a multiple of 8, explicit initial state) use the torch composite;
* CPU tensors  -> ATen's ``lstm``/``gru`` (the reference's own native path), used
  for the gloo plumbing configuration and as the numerical oracle in tests.
"""
from __future__ import annotations

import math
from typing import List, Optional, Tuple

import torch
from torch import nn

_VF = torch._VF  # ATen rnn entry points (CPU oracle path)


def rnn_composite(cell: str, x: torch.Tensor, weights: List[torch.Tensor], num_layers: int,
                  hidden: int, hx=None, batch_first: bool = True):
    """Plain-PyTorch restatement of the recurrence (matmul + pointwise, no cuDNN/ATen rnn).

    Independent oracle for the kernels and the path taken on CUDA for shapes outside the
    fused kernel families.  ``weights`` is [w_ih, w_hh, b_ih, b_hh] per layer.
    """
    if batch_first:
        x = x.transpose(0, 1)
    T, B, _ = x.shape
    h_n, c_n = [], []
    inp = x
    for l in range(num_layers):
        w_ih, w_hh, b_ih, b_hh = weights[4 * l:4 * l + 4]
        if hx is None:
            h = x.new_zeros(B, hidden)
            c = x.new_zeros(B, hidden)
        elif cell == "lstm":
            h, c = hx[0][l], hx[1][l]
        else:
            h, c = hx[l], None
        gx = torch.matmul(inp, w_ih.t()) + b_ih  # input projection hoisted over all timesteps
        outs = []
        for t in range(T):
            gh = torch.addmm(b_hh, h, w_hh.t())
            if cell == "lstm":
                i, f, g, o = (gx[t] + gh).chunk(4, dim=1)
                c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(g)
                h = torch.sigmoid(o) * torch.tanh(c)
            else:
                xr, xz, xn = gx[t].chunk(3, dim=1)
                hr, hz, hn = gh.chunk(3, dim=1)
                r = torch.sigmoid(xr + hr)
                z = torch.sigmoid(xz + hz)
                n = torch.tanh(xn + r * hn)
                h = (1 - z) * n + z * h
            outs.append(h)
        inp = torch.stack(outs, 0)
        h_n.append(h)
        if cell == "lstm":
            c_n.append(c)
    out = inp.transpose(0, 1) if batch_first else inp
    if cell == "lstm":
        return out, (torch.stack(h_n, 0), torch.stack(c_n, 0))
    return out, torch.stack(h_n, 0)


class FusedRNNBase(nn.Module):
    cell = "lstm"
    gates = 4

    def __init__(self, input_size: int, hidden_size: int, num_layers: int = 1, bias: bool = True,
                 batch_first: bool = False, dropout: float = 0.0, bidirectional: bool = False,
                 device=None, dtype=None):
        super().__init__()
        if not bias:
            raise NotImplementedError("bias=False is not supported")
        if bidirectional:
            raise NotImplementedError("bidirectional RNNs are not supported (the reference uses none)")
        if dropout != 0.0:
            raise NotImplementedError("inter-layer dropout is not supported (reference model has none, Q1)")
        if dtype not in (None, torch.float32):
            raise TypeError("master weights are fp32")
        self.input_size = input_size
        self.hidden_size = hidden_size
        self.num_layers = num_layers
        self.batch_first = batch_first
        G = self.gates
        for l in range(num_layers):
            in_l = input_size if l == 0 else hidden_size
            # creation order == nn.LSTM's, so state_dict order and RNG consumption match
            self.register_parameter(f"weight_ih_l{l}", nn.Parameter(torch.empty(G * hidden_size, in_l, device=device)))
            self.register_parameter(f"weight_hh_l{l}", nn.Parameter(torch.empty(G * hidden_size, hidden_size, device=device)))
            self.register_parameter(f"bias_ih_l{l}", nn.Parameter(torch.empty(G * hidden_size, device=device)))
            self.register_parameter(f"bias_hh_l{l}", nn.Parameter(torch.empty(G * hidden_size, device=device)))
        self.reset_parameters()

    def reset_parameters(self) -> None:
        stdv = 1.0 / math.sqrt(self.hidden_size) if self.hidden_size > 0 else 0
        for w in self.parameters():
            nn.init.uniform_(w, -stdv, stdv)

    def flat_weights(self) -> List[torch.Tensor]:
        out = []
        for l in range(self.num_layers):
            out += [getattr(self, f"weight_ih_l{l}"), getattr(self, f"weight_hh_l{l}"),
                    getattr(self, f"bias_ih_l{l}"), getattr(self, f"bias_hh_l{l}")]
        return out

    # -- nn.LSTM / nn.GRU API compatibility --
    def flatten_parameters(self) -> None:
        """No-op (kept so code written for ``nn.LSTM`` runs unchanged): the fused kernels read the weights in
        place — zero-copy when a :class:`~..utils.arena.ParamArena` owns the model — instead of cuDNN's packed blob."""

    @property
    def all_weights(self) -> List[List[torch.Tensor]]:
        """``[[weight_ih_l0, weight_hh_l0, bias_ih_l0, bias_hh_l0], ...]`` like ``nn.LSTM.all_weights``."""
        w = self.flat_weights()
        return [w[4 * l:4 * l + 4] for l in range(self.num_layers)]

    @property
    def bias(self) -> bool:
        return True

    @property
    def bidirectional(self) -> bool:
        return False

    @property
    def dropout(self) -> float:
        return 0.0

    @property
    def proj_size(self) -> int:
        return 0

    def extra_repr(self) -> str:
        return f"{self.input_size}, {self.hidden_size}, num_layers={self.num_layers}, batch_first={self.batch_first}"

    # -- dispatch --
    def forward(self, x: torch.Tensor, hx=None):
        if x.dim() != 3:
            raise ValueError("expected a 3-D input (batched sequences)")
        w = self.flat_weights()
        if x.is_cuda:
            from ..ops import rnn_function
            if hx is None and rnn_function.fused_supported(self, x):
                return rnn_function.fused_rnn(self, x)
            from ..ops import rnn_large
            if hx is None and rnn_large.large_supported(self, x):
                return rnn_large.large_rnn(self, x)
            return rnn_composite(self.cell, x, w, self.num_layers, self.hidden_size, hx, self.batch_first)
        return self._aten(x, w, hx)

    def _aten(self, x, w, hx):
        raise NotImplementedError


class FusedLSTM(FusedRNNBase):
    cell = "lstm"
    gates = 4

    def _aten(self, x, w, hx):
        B = x.size(0) if self.batch_first else x.size(1)
        if hx is None:
            z = x.new_zeros(self.num_layers, B, self.hidden_size)
            hx = (z, z)
        out, h, c = _VF.lstm(x, hx, w, True, self.num_layers, 0.0, self.training, False, self.batch_first)
        return out, (h, c)


class FusedGRU(FusedRNNBase):
    cell = "gru"
    gates = 3

    def _aten(self, x, w, hx):
        B = x.size(0) if self.batch_first else x.size(1)
        if hx is None:
            hx = x.new_zeros(self.num_layers, B, self.hidden_size)
        out, h = _VF.gru(x, hx, w, True, self.num_layers, 0.0, self.training, False, self.batch_first)
        return out, h
