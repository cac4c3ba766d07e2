"""Autograd glue between torch modules and the fused recurrent kernels.

``fused_rnn(module, x)`` behaves like ``nn.LSTM.forward`` (returns ``out, (h_n, c_n)``)
but runs ``b200rnn_{lstm,gru}_fwd`` / ``b200rnn_{lstm,gru}_bwd`` (csrc/rnn/lstm_small.inc) — the
replacement for ATen ``lstm`` + its autograd backward that the reference reaches through
``self.lstm(x)`` / ``loss.backward()`` (src/motion/model.py:14, trainer/base.py:116).
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Tuple

import torch

from .. import _native
from .._native import ModelDesc

CELL_ID = {"lstm": 0, "gru": 1}


def canonical_desc(cell: str, I: int, H: int, L: int, B: int, T: int, C_: int = 0) -> ModelDesc:
    """Descriptor for parameters packed contiguously in torch order (+ optional head)."""
    G = 4 if cell == "lstm" else 3
    d = ModelDesc()
    d.cell = CELL_ID[cell]
    d.input_size, d.hidden_size, d.num_layers = I, H, L
    d.batch, d.seq_len, d.num_classes = B, T, C_
    off = 0
    for l in range(L):
        in_l = I if l == 0 else H
        d.off_wih[l] = off; off += G * H * in_l
        d.off_whh[l] = off; off += G * H * H
        d.off_bih[l] = off; off += G * H
        d.off_bhh[l] = off; off += G * H
    if C_ > 0:
        d.off_fcw = off; off += C_ * H
        d.off_fcb = off; off += C_
    d.n_params = off
    return d


def packed_params(tensors: List[torch.Tensor]) -> torch.Tensor:
    """Return one contiguous fp32 vector holding ``tensors`` back to back.

    Zero-copy when they already are consecutive views of one arena (the normal case once a
    :class:`~..utils.arena.ParamArena` owns the model); otherwise a small ``torch.cat``.
    """
    base = tensors[0]
    expect = base.data_ptr()
    contiguous = True
    for t in tensors:
        if not t.is_contiguous() or t.data_ptr() != expect:
            contiguous = False
            break
        expect += t.numel() * 4
    total = sum(t.numel() for t in tensors)
    if contiguous:
        try:
            return base.detach().as_strided((total,), (1,), base.storage_offset())
        except RuntimeError:
            pass  # storage too small to cover the run: fall through to a copy
    return torch.cat([t.detach().reshape(-1) for t in tensors])


def fused_supported(module, x: torch.Tensor) -> bool:
    if not x.is_cuda or x.dtype != torch.float32:
        return False
    lib = _native.lib()  # on a CUDA machine the library is mandatory
    B, T = (x.size(0), x.size(1)) if module.batch_first else (x.size(1), x.size(0))
    d = canonical_desc(module.cell, module.input_size, module.hidden_size, module.num_layers, B, T, 0)
    return bool(lib.b200rnn_rnn_small_supported(C.byref(d)))


class _FusedRNNFn(torch.autograd.Function):
    """One persistent launch forward, one backward (+ fixed-order row reduction), LSTM or GRU."""

    @staticmethod
    def forward(ctx, x: torch.Tensor, cell: str, H: int, L: int, *weights: torch.Tensor):
        lib = _native.lib()
        B, T, I = x.shape
        x = x.contiguous()
        d = canonical_desc(cell, I, H, L, B, T, 0)
        params = packed_params(list(weights))
        stash = torch.empty(lib.b200rnn_rnn_stash_floats(C.byref(d)), dtype=torch.float32, device=x.device)
        fwd = lib.b200rnn_lstm_fwd if cell == "lstm" else lib.b200rnn_gru_fwd
        _native.check(fwd(C.byref(d), params.data_ptr(), x.data_ptr(), stash.data_ptr(), 1,
                          _native.stream_ptr(x.device)), cell + "_fwd")
        _native.count_launch()
        n = L * T * B * H
        h_seq = stash[:n].view(L, T, B, H)
        ctx.save_for_backward(x, params, stash)
        ctx.desc = d
        ctx.cell = cell
        ctx.shapes = [w.shape for w in weights]
        out = h_seq[L - 1].permute(1, 0, 2)  # [B,T,H] view of the time-major stash
        if cell == "lstm":
            c_seq = stash[n:2 * n].view(L, T, B, H)
            return out, h_seq[:, T - 1], c_seq[:, T - 1]
        return out, h_seq[:, T - 1]

    @staticmethod
    def backward(ctx, d_out, d_hn, d_cn=None):
        lib = _native.lib()
        x, params, stash = ctx.saved_tensors
        d = ctx.desc
        dev = x.device
        scratch = torch.empty(lib.b200rnn_rnn_bwd_scratch_floats(C.byref(d)), dtype=torch.float32, device=dev)
        rb, nc = C.c_int32(), C.c_int32()
        _native.check(lib.b200rnn_rnn_small_tiling(C.byref(d), C.byref(rb), C.byref(nc)), "tiling")
        d_out = d_out.contiguous() if d_out is not None else None
        d_hn = d_hn.contiguous() if d_hn is not None else None
        d_cn = d_cn.contiguous() if d_cn is not None else None
        sp = _native.stream_ptr(dev)
        head = (C.byref(d), params.data_ptr(), x.data_ptr(), stash.data_ptr(), None, 1.0,
                _native.ptr(d_out), _native.ptr(d_hn))
        if ctx.cell == "lstm":
            head = head + (_native.ptr(d_cn),)
        dx = None
        if ctx.needs_input_grad[0]:
            # the *_bwd_dx kernels also emit the input gradient, time-major and padded to H columns
            B, T, I = x.shape
            dx_pad = torch.empty(T, B, d.hidden_size, dtype=torch.float32, device=dev)
            fn = lib.b200rnn_lstm_bwd_dx if ctx.cell == "lstm" else lib.b200rnn_gru_bwd_dx
            rc = fn(*head, scratch.data_ptr(), dx_pad.data_ptr(), sp)
            dx = dx_pad[:, :, :I].permute(1, 0, 2)
        else:
            fn = lib.b200rnn_lstm_bwd if ctx.cell == "lstm" else lib.b200rnn_gru_bwd
            rc = fn(*head, scratch.data_ptr(), sp)
        _native.check(rc, ctx.cell + "_bwd")
        row = d.n_params + 2
        grad = torch.empty(row, dtype=torch.float32, device=dev)
        _native.check(lib.b200rnn_reduce_rows(scratch.data_ptr(), nc.value, row, grad.data_ptr(), 0, sp), "reduce_rows")
        _native.count_launch(2)
        grads, off = [], 0
        for shp in ctx.shapes:
            k = 1
            for s in shp:
                k *= s
            grads.append(grad[off:off + k].view(shp))
            off += k
        return (dx, None, None, None, *grads)


def fused_rnn(module, x: torch.Tensor):
    """``nn.LSTM`` / ``nn.GRU`` forward semantics on the fused small-hidden kernels."""
    xb = x if module.batch_first else x.transpose(0, 1)
    res = _FusedRNNFn.apply(xb, module.cell, module.hidden_size, module.num_layers, *module.flat_weights())
    out = res[0] if module.batch_first else res[0].transpose(0, 1)
    if module.cell == "lstm":
        return out, (res[1], res[2])
    return out, res[1]
