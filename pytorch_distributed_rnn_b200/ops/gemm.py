"""Python face of the tcgen05 GEMM (csrc/gemm/gemm_tc.cu): ``C (+)= A @ B.T + bias``.

A ``[M,K]`` and B ``[N,K]`` are bf16, row-major; C is fp32 or bf16.  This is the K1/K4 building
block: input projection over all timesteps, per-step recurrent projection at large hidden sizes
and weight-gradient GEMMs (see include/b200rnn.h) — the products ATen / MKL compute inside
``nn.LSTM`` and its autograd backward for the reference (src/motion/model.py:9-14, trainer/base.py:116).  Library GEMMs (cuBLAS via torch.matmul)
are used nowhere on these paths.
"""
from __future__ import annotations

from typing import Optional

import torch

from .. import _native


def gemm_bf16(a: torch.Tensor, b: torch.Tensor, bias: Optional[torch.Tensor] = None,
              out: Optional[torch.Tensor] = None, accumulate: bool = False,
              out_dtype: torch.dtype = torch.float32) -> torch.Tensor:
    assert a.dtype == torch.bfloat16 and b.dtype == torch.bfloat16 and a.is_cuda and b.is_cuda
    assert a.dim() == 2 and b.dim() == 2 and a.size(1) == b.size(1)
    assert a.stride(1) == 1 and b.stride(1) == 1, "operands must be K-contiguous"
    M, K = a.shape
    N = b.size(0)
    if out is None:
        assert not accumulate
        out = torch.empty(M, N, dtype=out_dtype, device=a.device)
    assert out.stride(1) == 1 and out.shape == (M, N) and out.dtype in (torch.float32, torch.bfloat16)
    lib = _native.lib()
    _native.check(lib.b200rnn_gemm_bf16(a.data_ptr(), b.data_ptr(), out.data_ptr(), _native.ptr(bias), M, N, K,
                                        a.stride(0), b.stride(0), out.stride(0), int(accumulate),
                                        int(out.dtype == torch.bfloat16), _native.stream_ptr(a.device)), "gemm_bf16")
    _native.count_launch()
    return out


def cast_bf16(x: torch.Tensor, transpose: bool = False) -> torch.Tensor:
    """fp32 -> bf16 (optionally transposing a 2-D tensor) with the library's own kernels."""
    lib = _native.lib()
    x = x.contiguous()
    if not transpose:
        out = torch.empty(x.shape, dtype=torch.bfloat16, device=x.device)
        _native.check(lib.b200rnn_cast_bf16(x.data_ptr(), out.data_ptr(), x.numel(), _native.stream_ptr(x.device)), "cast_bf16")
    else:
        R, C = x.shape
        out = torch.empty(C, R, dtype=torch.bfloat16, device=x.device)
        _native.check(lib.b200rnn_transpose_cast_bf16(x.data_ptr(), out.data_ptr(), R, C, C, R,
                                                      _native.stream_ptr(x.device)), "transpose_cast_bf16")
    _native.count_launch()
    return out


def gemm_status() -> int:
    import ctypes as C
    st = C.c_int32(0)
    _native.check(_native.lib().b200rnn_gemm_status(C.byref(st)), "gemm_status")
    return st.value


def persist_status() -> int:
    """Non-zero once a bounded wait inside the persistent recurrent kernels (csrc/rnn/rnn_persist.cu) timed out."""
    import ctypes as C
    st = C.c_int32(0)
    _native.check(_native.lib().b200rnn_persist_status(C.byref(st)), "persist_status")
    return st.value


def wave_status() -> int:
    """Non-zero once a bounded wait of the small-hidden layer-wavefront kernels (csrc/rnn/lstm_small.inc) timed out."""
    import ctypes as C
    st = C.c_int32(0)
    _native.check(_native.lib().b200rnn_wave_status(C.byref(st)), "wave_status")
    return st.value
