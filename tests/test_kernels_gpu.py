"""Numerics of the hand-written sm_100a kernels against plain PyTorch fp32 references."""
import copy
import ctypes as C

import pytest
import torch
from torch import nn

from pytorch_distributed_rnn_b200 import FusedAdam, FusedLSTM, MotionModel, ParamArena, _native
from pytorch_distributed_rnn_b200.ops.engine import FusedTrainEngine

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _stock_lstm(m: FusedLSTM) -> nn.LSTM:
    s = nn.LSTM(m.input_size, m.hidden_size, m.num_layers, batch_first=True)
    s.load_state_dict({k: v.detach().cpu() for k, v in m.state_dict().items()})
    return s


def test_native_library_is_loaded():
    assert _native.available(), "libb200rnn.so must be loadable on a GPU box"
    assert _native.lib().b200rnn_version() >= 100


@pytest.mark.parametrize("B,T,I,L", [(5, 7, 9, 1), (37, 16, 9, 2), (64, 128, 9, 2), (1440, 128, 9, 2), (33, 12, 20, 3), (300, 5, 32, 2)])
def test_lstm_forward_matches_torch(B, T, I, L):
    torch.manual_seed(B + T)
    m = FusedLSTM(I, 32, L, batch_first=True).to(DEV)
    ref = _stock_lstm(m)
    x = torch.randn(B, T, I)
    out, (h, c) = m(x.to(DEV))
    with torch.no_grad():
        ro, (rh, rc) = ref(x)
    assert out.shape == (B, T, 32) and h.shape == (L, B, 32)
    assert torch.allclose(out.cpu(), ro, atol=2e-5, rtol=1e-4), float((out.cpu() - ro).abs().max())
    assert torch.allclose(h.cpu(), rh, atol=2e-5, rtol=1e-4)
    assert torch.allclose(c.cpu(), rc, atol=2e-5, rtol=1e-4)


@pytest.mark.parametrize("B,T,I,L", [(6, 5, 9, 1), (37, 16, 9, 2), (130, 33, 9, 2), (19, 9, 24, 3)])
def test_lstm_backward_matches_torch(B, T, I, L):
    torch.manual_seed(7 * B + L)
    m = FusedLSTM(I, 32, L, batch_first=True).to(DEV)
    ref = _stock_lstm(m)
    x = torch.randn(B, T, I)
    wo, wh, wc = torch.randn(B, T, 32), torch.randn(L, B, 32), torch.randn(L, B, 32)
    out, (h, c) = m(x.to(DEV))
    ((out * wo.to(DEV)).sum() + (h * wh.to(DEV)).sum() + (c * wc.to(DEV)).sum()).backward()
    ro, (rh, rc) = ref(x)
    ((ro * wo).sum() + (rh * wh).sum() + (rc * wc).sum()).backward()
    for (n, p), q in zip(m.named_parameters(), ref.parameters()):
        err = float((p.grad.cpu() - q.grad).abs().max())
        scale = float(q.grad.abs().max()) + 1e-6
        assert err <= 2e-4 * scale + 2e-5, (n, err, scale)


def _ref_train(model_cpu, xs, ys, lr, steps):
    opt = torch.optim.Adam(model_cpu.parameters(), lr=lr)
    metrics = []
    for s in range(steps):
        opt.zero_grad()
        out = model_cpu(xs[s])
        loss = nn.functional.cross_entropy(out, ys[s])
        loss.backward()
        opt.step()
        metrics.append((loss.item(), int((out.argmax(1) == ys[s]).sum())))
    return metrics


@pytest.mark.parametrize("B,T,L", [(96, 32, 2), (1440, 128, 2), (50, 10, 1)])
def test_engine_train_steps_match_torch_adam(B, T, L):
    torch.manual_seed(B)
    steps, lr = 3, 0.0025
    model = MotionModel(9, 32, L, 6)
    ref = copy.deepcopy(model)
    xs = [torch.randn(B, T, 9) for _ in range(steps)]
    ys = [torch.randint(0, 6, (B,)) for _ in range(steps)]
    want = _ref_train(ref, xs, ys, lr, steps)
    model = model.to(DEV)
    eng = FusedTrainEngine(model, lr=lr)
    got = []
    for s in range(steps):
        m = eng.train_step(xs[s].to(DEV), ys[s].to(DEV).view(-1, 1)).cpu()
        got.append((float(m[0]) / B, int(round(float(m[1])))))
    for (gl, gc), (wl, wc) in zip(got, want):
        assert abs(gl - wl) < 2e-5, (got, want)
        assert abs(gc - wc) <= max(1, B // 500), (got, want)  # ties in argmax may flip on rounding
    for (n, p), q in zip(model.named_parameters(), ref.parameters()):
        assert torch.allclose(p.detach().cpu(), q.detach(), atol=3e-5), (n, float((p.detach().cpu() - q).abs().max()))
    # Adam state is exported in the stock format
    sd = eng.optimizer.state_dict()
    assert float(sd["state"][0]["step"]) == steps and sd["state"][0]["exp_avg"].shape == (128, 9)


@pytest.mark.parametrize("B,T,L,cell", [(180, 24, 2, "lstm"), (720, 8, 2, "lstm"), (90, 12, 3, "lstm"), (1440, 6, 2, "lstm"),
                                        (180, 16, 2, "gru")])
def test_engine_index_slices_and_wavefront_handover(B, T, L, cell):
    """The training-loop form of the step: batches are index slices of a device-resident dataset (read by the
    kernels' own loaders, no gather launch), and from the second step on no sentinel fill is launched because
    the backward put the sentinels back — 3 launches per step — with an evaluation in between (which leaves
    activations in the stash, so the step after it must fill again).  Oracle: torch fp32 + torch.optim.Adam."""
    torch.manual_seed(B + L)
    steps, lr, N = 5, 0.0025, 2 * B + 7
    model = MotionModel(9, 32, L, 6, cell=cell)
    ref = copy.deepcopy(model)
    feats, labels = torch.randn(N, T, 9), torch.randint(0, 6, (N, 1))
    idxs = [torch.randperm(N)[:B] for _ in range(steps)]
    want = _ref_train(ref, [feats[i] for i in idxs], [labels[i].view(-1) for i in idxs], lr, steps)
    model = model.to(DEV)
    eng = FusedTrainEngine(model, lr=lr)
    fd, ld = feats.to(DEV), labels.to(DEV)
    got, launches = [], []
    for s in range(steps):
        n0 = _native.launches()
        m = eng.train_step_gather(fd, ld, idxs[s].to(DEV)).cpu()
        launches.append(_native.launches() - n0)
        got.append((float(m[0]) / B, int(round(float(m[1])))))
        if s == 2:
            eng.evaluate(fd[:B].contiguous(), ld[:B].contiguous())
    _native.assert_device_healthy()
    for (gl, gc), (wl, wc) in zip(got, want):
        assert abs(gl - wl) < 2e-5, (got, want)
        assert abs(gc - wc) <= max(1, B // 500), (got, want)
    for (n, p), q in zip(model.named_parameters(), ref.parameters()):
        assert torch.allclose(p.detach().cpu(), q.detach(), atol=5e-5), (n, float((p.detach().cpu() - q).abs().max()))
    assert launches[1] == 3 and launches[2] == 3 and launches[4] == 3, launches
    assert launches[0] >= 3 and launches[3] >= 3, launches


def test_engine_evaluate_and_head():
    torch.manual_seed(0)
    model = MotionModel(9, 32, 2, 6)
    x, y = torch.randn(367, 128, 9), torch.randint(0, 6, (367, 1))
    with torch.no_grad():
        out = model(x)
        want_loss = nn.functional.cross_entropy(out, y.view(-1), reduction="sum").item()
        want_corr = int((out.argmax(1) == y.view(-1)).sum())
    eng = FusedTrainEngine(model.to(DEV))
    m = eng.evaluate(x.to(DEV), y.to(DEV)).cpu()
    assert abs(float(m[0]) - want_loss) < 2e-3 and abs(int(round(float(m[1]))) - want_corr) <= 1


def test_gather_kernel():
    lib = _native.lib()
    feats = torch.randn(500, 128, 9, device=DEV)
    labels = torch.randint(0, 6, (500, 1), device=DEV)
    idx = torch.randperm(500, device=DEV)[:77]
    xo = torch.empty(77, 128, 9, device=DEV)
    yo = torch.empty(77, dtype=torch.int64, device=DEV)
    _native.check(lib.b200rnn_gather_batch(feats.data_ptr(), labels.data_ptr(), idx.data_ptr(), 77, 128 * 9,
                                           xo.data_ptr(), yo.data_ptr(), _native.stream_ptr()))
    assert torch.equal(xo, feats[idx]) and torch.equal(yo, labels[idx].view(-1))


def test_fused_adam_kernel_matches_torch():
    torch.manual_seed(3)
    a = nn.Linear(37, 11).to(DEV)
    b = copy.deepcopy(a)
    fa = FusedAdam(ParamArena(a), lr=0.01)
    sa = torch.optim.Adam(b.parameters(), lr=0.01)
    for _ in range(4):
        x = torch.randn(9, 37, device=DEV)
        for mod, opt in ((a, fa), (b, sa)):
            opt.zero_grad()
            mod(x).pow(2).sum().backward()
            opt.step()
    for p, q in zip(a.parameters(), b.parameters()):
        assert torch.allclose(p, q, atol=1e-6)


def test_unsupported_shape_falls_back_to_composite_and_is_correct():
    torch.manual_seed(1)
    m = FusedLSTM(9, 44, 2, batch_first=True).to(DEV)  # H=44: outside both kernel families (not a multiple of 8)
    ref = nn.LSTM(9, 44, 2, batch_first=True)
    ref.load_state_dict({k: v.detach().cpu() for k, v in m.state_dict().items()})
    x = torch.randn(4, 6, 9)
    out, _ = m(x.to(DEV))
    assert torch.allclose(out.detach().cpu(), ref(x)[0], atol=1e-4)


@pytest.mark.parametrize("B,T,I,L", [(5, 7, 9, 1), (37, 16, 9, 2), (1440, 128, 9, 2), (33, 12, 20, 3), (300, 5, 32, 2)])
def test_tensor_core_forward_matches_torch_and_ffma(B, T, I, L):
    """tcgen05 recurrent kernel (3xTF32) against stock fp32 nn.LSTM and against the FFMA kernel's stash."""
    from pytorch_distributed_rnn_b200.ops.rnn_function import canonical_desc, packed_params
    lib = _native.lib()
    torch.manual_seed(B * 3 + T)
    m = FusedLSTM(I, 32, L, batch_first=True).to(DEV)
    ref = _stock_lstm(m)
    x = torch.randn(B, T, I)
    xd = x.to(DEV)
    d = canonical_desc("lstm", I, 32, L, B, T, 0)
    params = packed_params(m.flat_weights())
    n = lib.b200rnn_rnn_stash_floats(C.byref(d))
    s_tc = torch.zeros(n, device=DEV)
    s_ff = torch.zeros(n, device=DEV)
    sp = _native.stream_ptr()
    _native.check(lib.b200rnn_lstm_fwd_tc(C.byref(d), params.data_ptr(), xd.data_ptr(), s_tc.data_ptr(), 1, sp), "fwd_tc")
    _native.check(lib.b200rnn_lstm_fwd(C.byref(d), params.data_ptr(), xd.data_ptr(), s_ff.data_ptr(), 1, sp), "fwd")
    torch.cuda.synchronize()
    st = C.c_int32(-1)
    _native.check(lib.b200rnn_tc_status(C.byref(st)))
    assert st.value == 0, "tensor-core pipeline watchdog fired"
    with torch.no_grad():
        ro, _ = ref(x)
    out_tc = s_tc[:L * T * B * 32].view(L, T, B, 32)[L - 1].permute(1, 0, 2).cpu()
    assert torch.allclose(out_tc, ro, atol=3e-5, rtol=1e-4), float((out_tc - ro).abs().max())
    assert torch.allclose(s_tc, s_ff, atol=3e-5, rtol=1e-4), float((s_tc - s_ff).abs().max())


# ---- hidden size 16 (the other member of the small-hidden fused family; `--hidden-units 16`) ----
@pytest.mark.parametrize("cell,B,T,I,L", [("lstm", 37, 16, 9, 2), ("lstm", 1440, 12, 9, 2), ("lstm", 19, 9, 24, 3),
                                          ("gru", 50, 21, 9, 2), ("gru", 6, 5, 30, 1)])
def test_hidden16_forward_and_backward_match_torch(cell, B, T, I, L):
    from pytorch_distributed_rnn_b200 import FusedGRU
    torch.manual_seed(3 * B + L)
    m = (FusedLSTM if cell == "lstm" else FusedGRU)(I, 16, L, batch_first=True).to(DEV)
    ref = (nn.LSTM if cell == "lstm" else nn.GRU)(I, 16, L, batch_first=True)
    ref.load_state_dict({k: v.detach().cpu() for k, v in m.state_dict().items()})
    x = torch.randn(B, T, I)
    wo, wh = torch.randn(B, T, 16), torch.randn(L, B, 16)
    out, hx = m(x.to(DEV))
    ro, rhx = ref(x)
    h, rh = (hx[0], rhx[0]) if cell == "lstm" else (hx, rhx)
    assert torch.allclose(out.detach().cpu(), ro.detach(), atol=2e-5, rtol=1e-4), float((out.detach().cpu() - ro).abs().max())
    ((out * wo.to(DEV)).sum() + (h * wh.to(DEV)).sum()).backward()
    ((ro * wo).sum() + (rh * wh).sum()).backward()
    for (n, p), q in zip(m.named_parameters(), ref.parameters()):
        err = float((p.grad.cpu() - q.grad).abs().max())
        scale = float(q.grad.abs().max()) + 1e-6
        assert err <= 2e-4 * scale + 2e-5, (n, err, scale)
    _native.assert_device_healthy()


@pytest.mark.parametrize("B,T,L,cell", [(96, 32, 2, "lstm"), (180, 24, 2, "lstm"), (50, 10, 1, "gru")])
def test_engine_hidden16_train_steps_match_torch_adam(B, T, L, cell):
    torch.manual_seed(B + 16)
    steps, lr, N = 4, 0.0025, 2 * B + 3
    model = MotionModel(9, 16, L, 6, cell=cell)
    ref = copy.deepcopy(model)
    feats, labels = torch.randn(N, T, 9), torch.randint(0, 6, (N, 1))
    idxs = [torch.randperm(N)[:B] for _ in range(steps)]
    want = _ref_train(ref, [feats[i] for i in idxs], [labels[i].view(-1) for i in idxs], lr, steps)
    model = model.to(DEV)
    eng = FusedTrainEngine(model, lr=lr)                 # the 3-launch engine, not the composite fallback
    fd, ld = feats.to(DEV), labels.to(DEV)
    got = []
    for s in range(steps):
        m = eng.train_step_gather(fd, ld, idxs[s].to(DEV)).cpu()
        got.append((float(m[0]) / B, int(round(float(m[1])))))
    _native.assert_device_healthy()
    for (gl, gc), (wl, wc) in zip(got, want):
        assert abs(gl - wl) < 2e-5, (got, want)
        assert abs(gc - wc) <= max(1, B // 500), (got, want)
    for (n, p), q in zip(model.named_parameters(), ref.parameters()):
        assert torch.allclose(p.detach().cpu(), q.detach(), atol=5e-5), (n, float((p.detach().cpu() - q).abs().max()))
