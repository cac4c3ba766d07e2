"""The allocation-free backward, driven through the C ABI exactly as a non-Python host would.

``b200rnn_lstm_bwd_static`` exists for programs that cannot hand the backward pass a scratch
buffer — a C/C++ host without torch's allocator, a process whose own pool already owns the
whole GPU, a loop that cannot tolerate cudaMalloc's device-wide stall — so its scratch is an
array compiled into the library.  Here its result is checked against the regular
``b200rnn_lstm_bwd`` + ``b200rnn_reduce_rows`` pair on two calls in a row (the second with
different data, so contents left in the shared array by the first call must not leak through),
and an over-sized shape must be refused with a readable message.

This file sorts last on purpose: a run that stops at the first failure has then already run
every other test.  torch is used only to create inputs and read results.
"""
import ctypes as C

import pytest
import torch

from pytorch_distributed_rnn_b200 import _native
from pytorch_distributed_rnn_b200.ops.rnn_function import canonical_desc

pytestmark = pytest.mark.gpu

STATIC_SCRATCH_FLOATS = 4 * 1024 * 1024  # B200RNN_STATIC_SCRATCH_FLOATS in include/b200rnn.h


def _case(lib, seed, B, T, I, L, Cc, dev):
    g = torch.Generator().manual_seed(seed)
    d = canonical_desc("lstm", I, 32, L, B, T, Cc)
    params = ((torch.rand(d.n_params, generator=g) - 0.5) * 0.35).to(dev)
    x = torch.randn(B, T, I, generator=g).to(dev)
    labels = torch.randint(0, Cc, (B,), generator=g).to(dev)
    stash = torch.empty(lib.b200rnn_rnn_stash_floats(C.byref(d)), dtype=torch.float32, device=dev)
    sp = _native.stream_ptr()
    _native.check(lib.b200rnn_lstm_fwd(C.byref(d), params.data_ptr(), x.data_ptr(), stash.data_ptr(), 1, sp), "fwd")
    return d, params, x, labels, stash, sp


def _regular(lib, d, params, x, labels, stash, sp, dev):
    scratch = torch.empty(lib.b200rnn_rnn_bwd_scratch_floats(C.byref(d)), dtype=torch.float32, device=dev)
    rb, nc = C.c_int32(), C.c_int32()
    _native.check(lib.b200rnn_rnn_small_tiling(C.byref(d), C.byref(rb), C.byref(nc)))
    _native.check(lib.b200rnn_lstm_bwd(C.byref(d), params.data_ptr(), x.data_ptr(), stash.data_ptr(), labels.data_ptr(),
                                       1.0 / d.batch, None, None, None, scratch.data_ptr(), sp), "bwd")
    out = torch.empty(d.n_params + 2, dtype=torch.float32, device=dev)
    _native.check(lib.b200rnn_reduce_rows(scratch.data_ptr(), nc.value, d.n_params + 2, out.data_ptr(), 0, sp), "reduce")
    return out


def _static(lib, d, params, x, labels, stash, sp, dev):
    out = torch.full((d.n_params + 2,), float("nan"), dtype=torch.float32, device=dev)
    _native.check(lib.b200rnn_lstm_bwd_static(C.byref(d), params.data_ptr(), x.data_ptr(), stash.data_ptr(),
                                              labels.data_ptr(), 1.0 / d.batch, None, None, None, out.data_ptr(), sp),
                  "bwd_static")
    return out


def test_static_scratch_backward_matches_regular_entry_point_twice():
    lib = _native.lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    for seed, (B, T, I, L) in ((11, (96, 128, 9, 2)), (12, (50, 40, 20, 3))):
        case = _case(lib, seed, B, T, I, L, 6, dev)
        assert lib.b200rnn_rnn_bwd_scratch_floats(C.byref(case[0])) <= STATIC_SCRATCH_FLOATS
        want = _regular(lib, *case, dev)
        got = _static(lib, *case, dev)
        torch.cuda.synchronize()
        assert bool(torch.isfinite(got).all())
        scale = float(want[:-2].abs().max())
        assert torch.allclose(got, want, atol=2e-5 * max(scale, 1.0), rtol=1e-4), float((got - want).abs().max())
        assert float(got[-1]) == float(want[-1])  # correct count is an integer: must agree exactly


def test_static_scratch_backward_refuses_oversized_shape():
    lib = _native.lib()
    d = canonical_desc("lstm", 9, 32, 2, 1440, 128, 6)  # needs ~13.8 M floats, the array holds 4 Mi
    assert lib.b200rnn_rnn_bwd_scratch_floats(C.byref(d)) > STATIC_SCRATCH_FLOATS
    one = torch.zeros(8, device="cuda")  # never dereferenced: the size check comes first
    rc = lib.b200rnn_lstm_bwd_static(C.byref(d), one.data_ptr(), one.data_ptr(), one.data_ptr(), None, 1.0,
                                     None, None, None, one.data_ptr(), _native.stream_ptr())
    assert rc == -4  # B200RNN_ETOOBIG
    msg = lib.b200rnn_last_error().decode()
    assert "scratch" in msg and "b200rnn_lstm_bwd" in msg


def test_static_scratch_backward_with_output_gradients_three_layers():
    """Same comparison on the other gradient sources: an explicit gradient of the whole output sequence plus final-state
    gradients instead of the fused head, three layers, a batch small enough for the layer wavefront (the built-in
    scratch then also carries the inter-layer dx hand-off, sentinels included)."""
    lib = _native.lib()
    dev = torch.device("cuda", torch.cuda.current_device())
    B, T, I, L = 24, 33, 9, 3
    d, params, x, _, stash, sp = _case(lib, 21, B, T, I, L, 6, dev)
    g = torch.Generator().manual_seed(99)
    d_out = torch.randn(B, T, 32, generator=g).to(dev)
    d_hn = torch.randn(L, B, 32, generator=g).to(dev)
    d_cn = torch.randn(L, B, 32, generator=g).to(dev)
    assert lib.b200rnn_rnn_small_bwd_wave(C.byref(d)) == 1          # the shape does run as a wavefront
    scratch = torch.empty(lib.b200rnn_rnn_bwd_scratch_floats(C.byref(d)), dtype=torch.float32, device=dev)
    rb, nc = C.c_int32(), C.c_int32()
    _native.check(lib.b200rnn_rnn_small_tiling(C.byref(d), C.byref(rb), C.byref(nc)))
    outs = []
    for rep in range(2):     # twice: what the first call leaves in the shared array must not leak into the second
        _native.check(lib.b200rnn_lstm_bwd(C.byref(d), params.data_ptr(), x.data_ptr(), stash.data_ptr(), None, 1.0,
                                           d_out.data_ptr(), d_hn.data_ptr(), d_cn.data_ptr(), scratch.data_ptr(), sp), "bwd")
        want = torch.empty(d.n_params + 2, dtype=torch.float32, device=dev)
        _native.check(lib.b200rnn_reduce_rows(scratch.data_ptr(), nc.value, d.n_params + 2, want.data_ptr(), 0, sp), "reduce")
        got = torch.full((d.n_params + 2,), float("nan"), dtype=torch.float32, device=dev)
        _native.check(lib.b200rnn_lstm_bwd_static(C.byref(d), params.data_ptr(), x.data_ptr(), stash.data_ptr(), None, 1.0,
                                                  d_out.data_ptr(), d_hn.data_ptr(), d_cn.data_ptr(), got.data_ptr(), sp),
                      "bwd_static")
        torch.cuda.synchronize()
        scale = float(want[:-2].abs().max())
        assert bool(torch.isfinite(got).all())
        assert torch.allclose(got, want, atol=2e-5 * max(scale, 1.0), rtol=1e-4), float((got - want).abs().max())
        d_out = d_out * 0.5 + 1.0
        outs.append(got)
    assert not torch.equal(outs[0], outs[1])
    _native.assert_device_healthy()
