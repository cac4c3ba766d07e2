"""Out-of-bounds writes: every caller-owned buffer of the C ABI between two guard bands.

compute-sanitizer is not available on the GPU pool this project runs on ("Sanitizers"), so the bounds of
the kernels' writes are checked the slow way: each buffer handed to an entry point is carved out of a larger
allocation whose 16 KiB before and after it hold a known bit pattern; after the call both bands must be intact.
That covers what memcheck would report for the buffers the library does not own - in particular that the size
queries (`*_stash_floats`, `*_scratch_floats`, `*_bytes`) really bound what the kernels touch, for ragged shapes
(row counts and widths that are not multiples of any tile).  Read-only inputs are guarded as well: a kernel that
writes into its inputs (the wavefront re-arm does, on purpose, only with B200RNN_STEP_REARM) would show up.
Out-of-bounds READS stay undetected.
"""
import ctypes as C

import pytest
import torch

from pytorch_distributed_rnn_b200 import _native
from pytorch_distributed_rnn_b200.ops.rnn_function import canonical_desc

pytestmark = pytest.mark.gpu

GUARD = 4096          # elements of 4 bytes on each side
PATTERN = 0x5AFEC0DE  # as int32


class Guarded:
    """A device buffer of `nbytes` bytes with guard bands; .ptr is what the kernels get."""

    def __init__(self, nbytes, fill=None, dev="cuda"):
        words = (nbytes + 3) // 4
        self.raw = torch.full((words + 2 * GUARD,), PATTERN, dtype=torch.int32, device=dev)
        self.body = self.raw[GUARD:GUARD + words]
        self.nbytes = nbytes
        if fill is not None:
            self.f32.copy_(fill.reshape(-1).to(torch.float32))
        else:
            self.body.zero_()

    @property
    def f32(self):
        return self.body.view(torch.float32)

    @property
    def ptr(self):
        return self.body.data_ptr()

    def intact(self):
        lo, hi = self.raw[:GUARD], self.raw[GUARD + self.body.numel():]
        return bool((lo == PATTERN).all()) and bool((hi == PATTERN).all())


def _check(named):
    torch.cuda.synchronize()
    broken = [n for n, g in named.items() if not g.intact()]
    assert not broken, f"guard band overwritten next to: {broken}"
    _native.assert_device_healthy()


@pytest.mark.parametrize("cell,H,B,T,I,L", [
    ("lstm", 32, 97, 13, 9, 2),      # sequential layers, ragged last tile
    ("lstm", 32, 23, 17, 20, 3),     # layer wavefront (few rows), wide input
    ("gru", 32, 50, 9, 9, 2),
    ("lstm", 16, 61, 11, 9, 2),
    ("gru", 16, 19, 8, 5, 1),
])
def test_small_family_stays_inside_its_buffers(cell, H, B, T, I, L):
    lib = _native.lib()
    Cn = 6
    g = torch.Generator().manual_seed(B * T)
    d = canonical_desc(cell, I, H, L, B, T, Cn)
    assert lib.b200rnn_rnn_small_supported(C.byref(d)) == 1
    params = Guarded(4 * d.n_params, (torch.rand(d.n_params, generator=g) - 0.5) * 0.3)
    x = Guarded(4 * B * T * I, torch.randn(B, T, I, generator=g))
    labels_t = torch.randint(0, Cn, (B,), generator=g)
    labels = Guarded(8 * B)
    labels.body.view(torch.int64).copy_(labels_t)
    stash = Guarded(4 * lib.b200rnn_rnn_stash_floats(C.byref(d)))
    scratch = Guarded(4 * lib.b200rnn_rnn_bwd_scratch_floats(C.byref(d)))
    rb, nc = C.c_int32(), C.c_int32()
    _native.check(lib.b200rnn_rnn_small_tiling(C.byref(d), C.byref(rb), C.byref(nc)))
    out = Guarded(4 * (d.n_params + 2))
    sp = _native.stream_ptr()
    fwd = lib.b200rnn_lstm_fwd if cell == "lstm" else lib.b200rnn_gru_fwd
    named = dict(params=params, x=x, labels=labels, stash=stash, scratch=scratch, out=out)
    for _ in range(2):
        _native.check(fwd(C.byref(d), params.ptr, x.ptr, stash.ptr, 1, sp), "fwd")
        if cell == "lstm":
            _native.check(lib.b200rnn_lstm_bwd(C.byref(d), params.ptr, x.ptr, stash.ptr, labels.ptr, 1.0 / B, None, None, None,
                                               scratch.ptr, sp), "bwd")
        else:
            _native.check(lib.b200rnn_gru_bwd(C.byref(d), params.ptr, x.ptr, stash.ptr, labels.ptr, 1.0 / B, None, None,
                                              scratch.ptr, sp), "bwd")
        _native.check(lib.b200rnn_reduce_rows(scratch.ptr, nc.value, d.n_params + 2, out.ptr, 0, sp), "reduce_rows")
        _check(named)
    assert bool(torch.isfinite(out.f32).all()) and float(out.f32[:d.n_params].abs().sum()) > 0


def test_small_family_input_gradient_variant_stays_inside_its_buffers():
    lib = _native.lib()
    B, T, I, L, H = 45, 10, 12, 2, 32
    g = torch.Generator().manual_seed(7)
    d = canonical_desc("lstm", I, H, L, B, T, 0)
    params = Guarded(4 * d.n_params, (torch.rand(d.n_params, generator=g) - 0.5) * 0.3)
    x = Guarded(4 * B * T * I, torch.randn(B, T, I, generator=g))
    d_out = Guarded(4 * B * T * H, torch.randn(B, T, H, generator=g))
    stash = Guarded(4 * lib.b200rnn_rnn_stash_floats(C.byref(d)))
    scratch = Guarded(4 * lib.b200rnn_rnn_bwd_scratch_floats(C.byref(d)))
    dx = Guarded(4 * T * B * 32)
    sp = _native.stream_ptr()
    _native.check(lib.b200rnn_lstm_fwd(C.byref(d), params.ptr, x.ptr, stash.ptr, 1, sp), "fwd")
    _native.check(lib.b200rnn_lstm_bwd_dx(C.byref(d), params.ptr, x.ptr, stash.ptr, None, 1.0, d_out.ptr, None, None,
                                          scratch.ptr, dx.ptr, sp), "bwd_dx")
    _check(dict(params=params, x=x, d_out=d_out, stash=stash, scratch=scratch, dx=dx))
    assert float(dx.f32.abs().sum()) > 0


def test_gather_update_head_stay_inside_their_buffers():
    lib = _native.lib()
    sp = _native.stream_ptr()
    g = torch.Generator().manual_seed(3)
    # --- gathers: 37 rows of 128 x 9 out of 200
    N, B, T, I = 200, 37, 128, 9
    feats = Guarded(4 * N * T * I, torch.randn(N, T, I, generator=g))
    labs = Guarded(8 * N)
    labs.body.view(torch.int64).copy_(torch.randint(0, 6, (N,), generator=g))
    idx = Guarded(8 * B)
    idx.body.view(torch.int64).copy_(torch.randperm(N, generator=g)[:B])
    xo, yo, xtm = Guarded(4 * B * T * I), Guarded(8 * B), Guarded(4 * T * B * I)
    _native.check(lib.b200rnn_gather_batch(feats.ptr, labs.ptr, idx.ptr, B, T * I, xo.ptr, yo.ptr, sp), "gather")
    _native.check(lib.b200rnn_gather_batch_capped(feats.ptr, labs.ptr, idx.ptr, B, T * I, xo.ptr, yo.ptr, 5, sp), "gather_capped")
    _native.check(lib.b200rnn_gather_batch_tm(feats.ptr, idx.ptr, B, T, I, xtm.ptr, sp), "gather_tm")
    _check(dict(feats=feats, labs=labs, idx=idx, xo=xo, yo=yo, xtm=xtm))
    want = feats.f32.view(N, T, I)[idx.body.view(torch.int64)]
    assert torch.equal(xo.f32.view(B, T, I), want) and torch.equal(xtm.f32.view(T, B, I), want.transpose(0, 1))

    # --- fused optimizer: a parameter count that is no multiple of 4, several rows and one row
    from pytorch_distributed_rnn_b200._native import OptimDesc
    for n_rows, P in ((7, 14149), (1, 1003), (3, 5)):
        rows = Guarded(4 * n_rows * (P + 2), torch.randn(n_rows, P + 2, generator=g))
        p, m, v, go, met = (Guarded(4 * P, torch.randn(P, generator=g)), Guarded(4 * P), Guarded(4 * P), Guarded(4 * P),
                            Guarded(8))
        od = OptimDesc()
        od.kind, od.step, od.lr, od.beta1, od.beta2, od.eps, od.grad_scale = 0, 1, 1e-3, 0.9, 0.999, 1e-8, 1.0
        _native.check(lib.b200rnn_fused_update(rows.ptr, n_rows, P, p.ptr, m.ptr, v.ptr, C.byref(od), go.ptr, met.ptr, sp),
                      "fused_update")
        _check(dict(rows=rows, p=p, m=m, v=v, grad_out=go, metrics=met))
        assert torch.allclose(go.f32, rows.f32.view(n_rows, P + 2)[:, :P].sum(0), atol=1e-5)

    # --- head + cross-entropy, ragged hidden and batch
    for B2, H2, C2 in ((33, 40, 6), (64, 1024, 6), (5, 8, 3)):
        h = Guarded(4 * B2 * H2, torch.randn(B2, H2, generator=g))
        w, b = Guarded(4 * C2 * H2, torch.randn(C2, H2, generator=g) * 0.1), Guarded(4 * C2, torch.randn(C2, generator=g))
        y = Guarded(8 * B2)
        y.body.view(torch.int64).copy_(torch.randint(0, C2, (B2,), generator=g))
        dh, dw, db, met = Guarded(4 * B2 * H2), Guarded(4 * C2 * H2), Guarded(4 * C2), Guarded(8)
        sc = Guarded(4 * lib.b200rnn_head_ce_scratch_floats(B2, C2))
        _native.check(lib.b200rnn_head_ce(h.ptr, H2, w.ptr, b.ptr, y.ptr, None, B2, H2, C2, 1.0 / B2, dh.ptr, H2, dw.ptr,
                                          db.ptr, met.ptr, sc.ptr, None, None, sp), "head_ce")
        logits = Guarded(4 * B2 * C2)
        # the forward-only head keeps its per-block partial sums behind metrics[0..1] (found by this very test)
        met2 = Guarded(4 * lib.b200rnn_head_fwd_metrics_floats(B2))
        _native.check(lib.b200rnn_head_fwd(h.ptr, H2, w.ptr, b.ptr, y.ptr, B2, H2, C2, logits.ptr, met2.ptr, sp), "head_fwd")
        _check(dict(h=h, w=w, b=b, y=y, dh=dh, dw=dw, db=db, metrics=met, scratch=sc, logits=logits, metrics_fwd=met2))
        assert torch.allclose(met2.f32[:2], met.f32, rtol=1e-5)      # both forms agree on (loss sum, correct count)


@pytest.mark.parametrize("M,N,K,bf16_out", [
    (200, 136, 72, False),      # ragged everything: one-tile-per-CTA kernel, partial tiles on both edges
    (130, 264, 40, True),
    (512, 512, 256, False),     # persistent CTA-pair kernel (TMA-store epilogue)
    (384, 768, 128, True),
    (128, 1024, 64, False),     # persistent single-CTA kernel (M < 256)
    (8, 4096, 1024, False),     # skinny M (the per-step recurrent product)
])
def test_gemm_stays_inside_its_output(M, N, K, bf16_out):
    lib = _native.lib()
    g = torch.Generator().manual_seed(M + N + K)
    a_t, b_t = torch.randn(M, K, generator=g).to(torch.bfloat16), torch.randn(N, K, generator=g).to(torch.bfloat16)
    a, b = Guarded(2 * M * K), Guarded(2 * N * K)
    a.raw[GUARD:GUARD + a.body.numel()].view(torch.bfloat16)[:M * K].copy_(a_t.reshape(-1))
    b.raw[GUARD:GUARD + b.body.numel()].view(torch.bfloat16)[:N * K].copy_(b_t.reshape(-1))
    bias = Guarded(4 * N, torch.randn(N, generator=g))
    c = Guarded((2 if bf16_out else 4) * M * N)
    _native.check(lib.b200rnn_gemm_bf16(a.ptr, b.ptr, c.ptr, bias.ptr, M, N, K, K, K, N, 0, int(bf16_out),
                                        _native.stream_ptr()), "gemm")
    _check(dict(a=a, b=b, bias=bias, c=c))
    got = (c.body.view(torch.bfloat16)[:M * N].float() if bf16_out else c.f32).view(M, N)
    want = a_t.float().cuda() @ b_t.float().cuda().t() + bias.f32
    assert float((got - want).abs().max()) < (0.5 if bf16_out else 2e-2) * max(1.0, float(want.abs().max()) / 32)
    # transposing cast, ragged
    src = Guarded(4 * M * K, torch.randn(M, K, generator=g))
    dst = Guarded(2 * K * M)
    _native.check(lib.b200rnn_transpose_cast_bf16(src.ptr, dst.ptr, M, K, K, M, _native.stream_ptr()), "transpose_cast")
    _check(dict(src=src, dst=dst))


@pytest.mark.parametrize("cell,I,H,L,B,T", [
    ("lstm", 9, 64, 2, 23, 11),
    ("lstm", 9, 1024, 1, 64, 6),     # persistent kernels at the scaled geometry
    ("gru", 24, 256, 2, 80, 5),      # clustered 2-D backward
    ("lstm", 20, 192, 1, 130, 4),    # per-timestep formulation
    ("gru", 9, 40, 2, 16, 7),
])
def test_large_path_stays_inside_its_buffers(cell, I, H, L, B, T):
    lib = _native.lib()
    g = torch.Generator().manual_seed(H + B)
    d = canonical_desc(cell, I, H, L, B, T, 0)
    params = Guarded(4 * d.n_params, (torch.rand(d.n_params, generator=g) - 0.5) * 0.2)
    x = Guarded(4 * T * B * I, torch.randn(T, B, I, generator=g))
    stash = Guarded(lib.b200rnn_rnn_large_stash_bytes(C.byref(d)))
    scratch = Guarded(lib.b200rnn_rnn_large_scratch_bytes(C.byref(d)))
    d_out = Guarded(4 * T * B * H, torch.randn(T, B, H, generator=g))
    d_hn = Guarded(4 * L * B * H, torch.randn(L, B, H, generator=g))
    grad, dx = Guarded(4 * d.n_params), Guarded(4 * T * B * I)
    sp = _native.stream_ptr()
    named = dict(params=params, x=x, stash=stash, scratch=scratch, d_out=d_out, d_hn=d_hn, grad=grad, dx=dx)
    for _ in range(2):
        _native.check(lib.b200rnn_rnn_large_fwd(C.byref(d), params.ptr, x.ptr, stash.ptr, scratch.ptr, sp), "large_fwd")
        _native.check(lib.b200rnn_rnn_large_bwd(C.byref(d), params.ptr, x.ptr, stash.ptr, scratch.ptr, d_out.ptr, d_hn.ptr,
                                                None, grad.ptr, dx.ptr, sp), "large_bwd")
        _check(named)
    assert bool(torch.isfinite(grad.f32).all()) and float(grad.f32.abs().sum()) > 0 and float(dx.f32.abs().sum()) > 0
