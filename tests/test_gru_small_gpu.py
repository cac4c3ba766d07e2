"""Fused small-hidden GRU kernels and the d/dx backward variants against stock torch fp32 (CPU)."""
import copy

import pytest
import torch
from torch import nn

from pytorch_distributed_rnn_b200 import FusedGRU, FusedLSTM, MotionModel, _native
from pytorch_distributed_rnn_b200.ops import rnn_function
from pytorch_distributed_rnn_b200.ops.engine import FusedTrainEngine

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _stock(m):
    cls = nn.GRU if m.cell == "gru" else nn.LSTM
    s = cls(m.input_size, m.hidden_size, m.num_layers, batch_first=m.batch_first)
    s.load_state_dict({k: v.detach().cpu() for k, v in m.state_dict().items()})
    return s


def _assert_grads_close(m, ref, tol=2e-4):
    for (n, p), q in zip(m.named_parameters(), ref.parameters()):
        assert p.grad is not None, n
        err = float((p.grad.cpu() - q.grad).abs().max())
        scale = float(q.grad.abs().max()) + 1e-6
        assert err <= tol * scale + 2e-5, (n, err, scale)


def test_gru_takes_the_fused_path():
    m = FusedGRU(9, 32, 2, batch_first=True).to(DEV)
    x = torch.randn(8, 4, 9, device=DEV)
    assert rnn_function.fused_supported(m, x)
    before = _native.launches()
    m(x)
    assert _native.launches() == before + 1   # one persistent launch for the whole sequence


@pytest.mark.parametrize("B,T,I,L", [(5, 7, 9, 1), (37, 16, 9, 2), (64, 128, 9, 2), (1440, 128, 9, 2), (33, 12, 20, 3), (300, 5, 32, 2)])
def test_gru_forward_matches_torch(B, T, I, L):
    torch.manual_seed(B + T)
    m = FusedGRU(I, 32, L, batch_first=True).to(DEV)
    ref = _stock(m)
    x = torch.randn(B, T, I)
    out, h = m(x.to(DEV))
    with torch.no_grad():
        ro, rh = ref(x)
    assert out.shape == (B, T, 32) and h.shape == (L, B, 32)
    assert torch.allclose(out.cpu(), ro, atol=2e-5, rtol=1e-4), float((out.cpu() - ro).abs().max())
    assert torch.allclose(h.cpu(), rh, atol=2e-5, rtol=1e-4)


@pytest.mark.parametrize("B,T,I,L", [(6, 5, 9, 1), (37, 16, 9, 2), (130, 33, 9, 2), (19, 9, 24, 3), (1440, 24, 9, 2)])
def test_gru_backward_matches_torch(B, T, I, L):
    torch.manual_seed(7 * B + L)
    m = FusedGRU(I, 32, L, batch_first=True).to(DEV)
    ref = _stock(m)
    x = torch.randn(B, T, I)
    wo, wh = torch.randn(B, T, 32), torch.randn(L, B, 32)
    out, h = m(x.to(DEV))
    ((out * wo.to(DEV)).sum() + (h * wh.to(DEV)).sum()).backward()
    ro, rh = ref(x)
    ((ro * wo).sum() + (rh * wh).sum()).backward()
    _assert_grads_close(m, ref)


def test_gru_time_major_layout():
    torch.manual_seed(3)
    m = FusedGRU(9, 32, 2, batch_first=False).to(DEV)
    ref = _stock(m)
    x = torch.randn(11, 6, 9)   # [T, B, I]
    out, h = m(x.to(DEV))
    ro, rh = ref(x)
    assert torch.allclose(out.detach().cpu(), ro, atol=2e-5, rtol=1e-4)
    assert torch.allclose(h.detach().cpu(), rh, atol=2e-5, rtol=1e-4)


@pytest.mark.parametrize("cell", ["lstm", "gru"])
@pytest.mark.parametrize("B,T,I,L", [(6, 5, 9, 1), (37, 16, 9, 2), (130, 12, 32, 2), (21, 9, 20, 3)])
def test_input_gradient_matches_torch(cell, B, T, I, L):
    """x.requires_grad routes backward through the *_bwd_dx kernels: dx and all dW must match."""
    torch.manual_seed(11 * B + L)
    m = (FusedGRU if cell == "gru" else FusedLSTM)(I, 32, L, batch_first=True).to(DEV)
    ref = _stock(m)
    x = torch.randn(B, T, I)
    xg = x.to(DEV).requires_grad_(True)
    xr = x.clone().requires_grad_(True)
    wo = torch.randn(B, T, 32)
    before = _native.launches()
    out = m(xg)[0]
    (out * wo.to(DEV)).sum().backward()
    assert _native.launches() == before + 3, "fused forward + backward(+dx) + row reduction expected"
    (ref(xr)[0] * wo).sum().backward()
    assert xg.grad is not None and xg.grad.shape == (B, T, I)
    err = float((xg.grad.cpu() - xr.grad).abs().max())
    scale = float(xr.grad.abs().max()) + 1e-6
    assert err <= 2e-4 * scale + 2e-5, (err, scale)
    _assert_grads_close(m, ref)


def test_stacked_under_a_trainable_layer():
    """The use case for dx: the recurrent layer sits on top of another trainable layer."""
    torch.manual_seed(5)
    emb, rnn = nn.Linear(12, 9), FusedGRU(9, 32, 2, batch_first=True)
    emb_r, rnn_r = copy.deepcopy(emb), _stock(rnn)
    emb, rnn = emb.to(DEV), rnn.to(DEV)
    x = torch.randn(40, 10, 12)
    rnn(emb(x.to(DEV)))[0][:, -1].square().sum().backward()
    rnn_r(emb_r(x))[0][:, -1].square().sum().backward()
    assert torch.allclose(emb.weight.grad.cpu(), emb_r.weight.grad, atol=1e-4, rtol=1e-3)
    assert torch.allclose(emb.bias.grad.cpu(), emb_r.bias.grad, atol=1e-4, rtol=1e-3)


def _ref_train(model_cpu, xs, ys, lr):
    opt = torch.optim.Adam(model_cpu.parameters(), lr=lr)
    metrics = []
    for x, y in zip(xs, ys):
        opt.zero_grad()
        out = model_cpu(x)
        loss = nn.functional.cross_entropy(out, y)
        loss.backward()
        opt.step()
        metrics.append((loss.item(), int((out.argmax(1) == y).sum())))
    return metrics


@pytest.mark.parametrize("B,T,L", [(96, 32, 2), (1440, 128, 2), (50, 10, 1)])
def test_engine_trains_a_gru_motion_model(B, T, L):
    """The 3-launch engine (fwd, head+CE+BPTT, reduce+Adam) on the GRU cell vs torch autograd + Adam."""
    torch.manual_seed(B)
    steps, lr = 3, 0.0025
    model = MotionModel(9, 32, L, 6, cell="gru")
    ref = nn.Module()
    ref.lstm, ref.fc = _stock(model.lstm), copy.deepcopy(model.fc)
    ref.forward = lambda x: ref.fc(ref.lstm(x)[0][:, -1, :])
    xs = [torch.randn(B, T, 9) for _ in range(steps)]
    ys = [torch.randint(0, 6, (B,)) for _ in range(steps)]
    want = _ref_train(ref, xs, ys, lr)
    model = model.to(DEV)
    eng = FusedTrainEngine(model, lr=lr)
    for s in range(steps):
        m = eng.train_step(xs[s].to(DEV), ys[s].to(DEV).view(-1, 1)).cpu()
        assert abs(float(m[0]) / B - want[s][0]) < 2e-5, (s, float(m[0]) / B, want[s])
        assert abs(int(round(float(m[1]))) - want[s][1]) <= max(1, B // 500)
    for (n, p), q in zip(model.named_parameters(), list(ref.lstm.parameters()) + list(ref.fc.parameters())):
        assert torch.allclose(p.detach().cpu(), q.detach(), atol=3e-5), (n, float((p.detach().cpu() - q).abs().max()))
    # evaluation path (forward-only kernel + head)
    with torch.no_grad():
        out = ref(xs[0])
        want_loss = nn.functional.cross_entropy(out, ys[0], reduction="sum").item()
    got = eng.evaluate(xs[0].to(DEV), ys[0].to(DEV)).cpu()
    assert abs(float(got[0]) - want_loss) < 2e-3 * max(1.0, want_loss / 100)


@pytest.mark.parametrize("cell", ["lstm", "gru"])
@pytest.mark.parametrize("B,T,L", [(180, 128, 2), (90, 40, 3), (360, 64, 2)])
def test_layer_wavefront_shapes(cell, B, T, L):
    """Per-rank batches of the 8 / 4-GPU strong-scaling runs: few tiles, so the layers of a tile run concurrently on
    different SMs (sentinel hand-off through global memory).  Forward, every gradient and the watchdog flag."""
    from pytorch_distributed_rnn_b200.ops.gemm import wave_status
    torch.manual_seed(B + L)
    m = (FusedGRU if cell == "gru" else FusedLSTM)(9, 32, L, batch_first=True).to(DEV)
    ref = _stock(m)
    x = torch.randn(B, T, 9)
    wo = torch.randn(B, T, 32)
    for _ in range(2):   # twice: the second call reuses cached allocations holding stale (non-sentinel) values
        m.zero_grad()
        out = m(x.to(DEV))[0]
        (out * wo.to(DEV)).sum().backward()
    ro = ref(x)[0]
    (ro * wo).sum().backward()
    assert torch.allclose(out.detach().cpu(), ro.detach(), atol=2e-5, rtol=1e-4)
    _assert_grads_close(m, ref)
    assert wave_status() == 0
