"""Autograd glue for the large-hidden LSTM / GRU path (csrc/rnn/rnn_large.cu): what ``nn.LSTM(..., batch_first=True)``
does for the reference model when ``--hidden-units`` is large (src/motion/model.py:9-14, main.py:21).

Covers every shape with ``hidden % 8 == 0`` that the small fused family does not (in particular the
scaled configuration of BASELINE.json: 2 layers, hidden 1024, seq_len 512, bf16): the input
projection over all timesteps, the per-step recurrent projection and the weight-gradient products
are tcgen05 GEMMs; gates + cell update are one fused pointwise kernel per step; the time loop runs
natively.  bf16 operands, fp32 accumulation and state; parameters stay fp32 (master weights) and
gradients come back fp32 in torch's parameter layout.
"""
from __future__ import annotations

import ctypes as C

import torch

from .. import _native
from .rnn_function import canonical_desc, packed_params


def large_supported(module, x: torch.Tensor) -> bool:
    return x.is_cuda and x.dtype == torch.float32 and module.hidden_size % 8 == 0 and module.hidden_size >= 8


class _LargeRNNFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x_tm: torch.Tensor, cell: str, H: int, L: int, *weights: torch.Tensor):
        lib = _native.lib()
        T, B, I = x_tm.shape
        x_tm = x_tm.contiguous()
        d = canonical_desc(cell, I, H, L, B, T, 0)
        params = packed_params(list(weights))
        dev = x_tm.device
        stash = torch.empty(lib.b200rnn_rnn_large_stash_bytes(C.byref(d)), dtype=torch.uint8, device=dev)
        scratch = torch.empty(lib.b200rnn_rnn_large_scratch_bytes(C.byref(d)), dtype=torch.uint8, device=dev)
        _native.check(lib.b200rnn_rnn_large_fwd(C.byref(d), params.data_ptr(), x_tm.data_ptr(), stash.data_ptr(),
                                                scratch.data_ptr(), _native.stream_ptr(dev)), "rnn_large_fwd")
        _native.count_launch(T * L * 2)

        def seq(off):
            return stash[off:off + T * B * H * 4].view(torch.float32).view(T, B, H)
        hs = [seq(lib.b200rnn_rnn_large_h_offset(C.byref(d), l)) for l in range(L)]
        out = hs[L - 1]
        h_n = torch.stack([h[T - 1] for h in hs], 0)
        ctx.save_for_backward(x_tm, params, stash)
        ctx.desc, ctx.cell, ctx.scratch = d, cell, scratch
        ctx.shapes = [w.shape for w in weights]
        if cell == "lstm":
            c_n = torch.stack([seq(lib.b200rnn_rnn_large_c_offset(C.byref(d), l))[T - 1] for l in range(L)], 0)
            return out, h_n, c_n
        return out, h_n

    @staticmethod
    def backward(ctx, d_out, d_hn, d_cn=None):
        lib = _native.lib()
        x_tm, params, stash = ctx.saved_tensors
        d = ctx.desc
        dev = x_tm.device
        grad = torch.zeros(d.n_params, dtype=torch.float32, device=dev)
        d_out = d_out.contiguous() if d_out is not None else None
        d_hn = d_hn.contiguous() if d_hn is not None else None
        d_cn = d_cn.contiguous() if d_cn is not None else None
        dx = torch.empty_like(x_tm) if ctx.needs_input_grad[0] else None
        _native.check(lib.b200rnn_rnn_large_bwd(C.byref(d), params.data_ptr(), x_tm.data_ptr(), stash.data_ptr(),
                                                ctx.scratch.data_ptr(), _native.ptr(d_out), _native.ptr(d_hn),
                                                _native.ptr(d_cn), grad.data_ptr(), _native.ptr(dx),
                                                _native.stream_ptr(dev)), "rnn_large_bwd")
        _native.count_launch(d.seq_len * d.num_layers * 2)
        grads, off = [], 0
        for shp in ctx.shapes:
            k = 1
            for s in shp:
                k *= s
            grads.append(grad[off:off + k].view(shp))
            off += k
        return (dx, None, None, None, *grads)


def large_rnn(module, x: torch.Tensor):
    x_tm = x.transpose(0, 1) if module.batch_first else x
    res = _LargeRNNFn.apply(x_tm, module.cell, module.hidden_size, module.num_layers, *module.flat_weights())
    out = res[0].transpose(0, 1) if module.batch_first else res[0]
    if module.cell == "lstm":
        return out, (res[1], res[2])
    return out, res[1]
