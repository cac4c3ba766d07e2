"""Host-side tiling / layer-wavefront planning of the small-hidden family (csrc/rnn/lstm_small.inc).

The planners are plain host code of the native library (they assume 148 SMs when no GPU is present), so the
decisions the 1 / 2 / 4 / 8-GPU strong-scaling runs depend on can be pinned without a GPU."""
import ctypes as C

import pytest

from pytorch_distributed_rnn_b200 import _native
from pytorch_distributed_rnn_b200.ops.rnn_function import canonical_desc

pytestmark = pytest.mark.skipif(not _native.available(), reason="native library not built")


def _plan(B, L=2, cell="lstm"):
    lib = _native.lib()
    d = canonical_desc(cell, 9, 32, L, B, 128, 6)
    rb, nc = C.c_int32(), C.c_int32()
    _native.check(lib.b200rnn_rnn_small_tiling(C.byref(d), C.byref(rb), C.byref(nc)), "tiling")
    return dict(rows=rb.value, tiles=nc.value, fwd_wave=bool(lib.b200rnn_rnn_small_fwd_wave(C.byref(d))),
                bwd_wave=bool(lib.b200rnn_rnn_small_bwd_wave(C.byref(d))),
                scratch=lib.b200rnn_rnn_bwd_scratch_floats(C.byref(d)), n_params=d.n_params)


def test_full_batch_keeps_layers_in_sequence():
    p = _plan(1440)
    assert not p["fwd_wave"] and not p["bwd_wave"]
    assert p["rows"] == 10 and p["tiles"] == 144          # one backward CTA per SM, a single wave of CTAs


@pytest.mark.parametrize("B", [180, 360])
def test_small_per_rank_batches_run_as_a_layer_wavefront(B):
    p = _plan(B)
    assert p["fwd_wave"] and p["bwd_wave"]
    assert p["rows"] <= 6                                  # backward wave kernels exist for tiles of 1..6 rows
    assert p["tiles"] * 2 <= 148                           # every (layer, tile) CTA resident at once
    assert p["tiles"] * p["rows"] >= B


def test_tiles_cover_the_batch_and_scratch_holds_rows_plus_dx():
    for B in (1, 7, 96, 180, 720, 1440, 6912):
        p = _plan(B)
        assert p["tiles"] * p["rows"] >= B > (p["tiles"] - 1) * p["rows"]
        assert p["scratch"] == p["tiles"] * (p["n_params"] + 2) + 2 * 128 * B * 32


def test_single_layer_and_deep_stacks_never_wave_in_backward():
    assert not _plan(180, L=1)["bwd_wave"] and not _plan(180, L=1)["fwd_wave"]
    assert not _plan(64, L=4)["bwd_wave"]                  # dx buffers ping-pong by layer parity: <= 3 concurrent layers


def test_gru_uses_the_same_plan():
    assert _plan(180, cell="gru") == {**_plan(180, cell="lstm"), "n_params": _plan(180, cell="gru")["n_params"],
                                      "scratch": _plan(180, cell="gru")["scratch"]}


# ---- reducer bucket plan: native (csrc/runtime/reducer.cpp) vs the Python planner, arbitrary parameter lists ----
from hypothesis import given, settings, strategies as st  # noqa: E402


@settings(max_examples=80, deadline=None)
@given(sizes=st.lists(st.integers(1, 3_000_000), min_size=1, max_size=40),
       first_kb=st.integers(1, 2048), cap_kb=st.integers(1, 32768))
def test_bucket_plan_native_equals_python_and_respects_caps(sizes, first_kb, cap_kb):
    from pytorch_distributed_rnn_b200.parallel.ddp import plan_buckets
    lib = _native.lib()
    nbytes = [4 * s for s in sizes]
    first, cap = first_kb << 10, cap_kb << 10
    arr = (C.c_int64 * len(nbytes))(*nbytes)
    out = (C.c_int32 * len(nbytes))()
    nb = C.c_int32()
    assert lib.b200rnn_bucket_plan(arr, len(nbytes), first, cap, out, C.byref(nb)) == 0
    py = plan_buckets(nbytes, first, cap)
    assert nb.value == len(py)
    for b, ids in enumerate(py):
        assert all(out[i] == b for i in ids)
        # a bucket only exceeds its cap when a single parameter does
        limit = first if b == 0 else cap
        assert sum(nbytes[i] for i in ids) <= limit or len(ids) == 1
    # reverse parameter order, every parameter exactly once
    flat = [i for ids in py for i in ids]
    assert flat == list(reversed(range(len(nbytes))))
