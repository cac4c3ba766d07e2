"""Large-hidden training engine (per-layer gradient buckets) and the standalone head + CE kernels against
plain PyTorch fp32 references."""
import copy
import ctypes as C

import pytest
import torch
from torch import nn

from pytorch_distributed_rnn_b200 import MotionModel, _native
from pytorch_distributed_rnn_b200.ops.layered_engine import LayeredTrainEngine

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.mark.parametrize("B,H,Cn", [(7, 40, 6), (64, 1024, 6), (130, 96, 31)])
def test_head_ce_kernels_match_torch(B, H, Cn):
    torch.manual_seed(B)
    lib = _native.lib()
    h = torch.randn(B, H, requires_grad=True)
    fc = nn.Linear(H, Cn)
    N = B + 5
    labels = torch.randint(0, Cn, (N,))
    idx = torch.randperm(N)[:B]
    logits = fc(h)
    loss = nn.functional.cross_entropy(logits, labels[idx])
    loss.backward()
    want_corr = int((logits.argmax(1) == labels[idx]).sum())
    hd, wd, bd = h.detach().to(DEV), fc.weight.detach().to(DEV), fc.bias.detach().to(DEV)
    d_h = torch.empty(B, H, device=DEV)
    d_w, d_b, met = torch.empty(Cn, H, device=DEV), torch.empty(Cn, device=DEV), torch.zeros(2, device=DEV)
    scratch = torch.empty(lib.b200rnn_head_ce_scratch_floats(B, Cn), device=DEV)
    counter = torch.zeros(1, dtype=torch.int32, device=DEV)
    cnt = C.c_int32(0)
    labels_d, idx_d = labels.to(DEV), idx.to(DEV)      # named: a temporary would be freed before the kernel reads it
    _native.check(lib.b200rnn_head_ce(hd.data_ptr(), H, wd.data_ptr(), bd.data_ptr(), labels_d.data_ptr(),
                                      idx_d.data_ptr(), B, H, Cn, 1.0 / B, d_h.data_ptr(), H, d_w.data_ptr(),
                                      d_b.data_ptr(), met.data_ptr(), scratch.data_ptr(), counter.data_ptr(), C.byref(cnt),
                                      _native.stream_ptr()), "head_ce")
    torch.cuda.synchronize()
    assert int(counter.item()) == cnt.value > 0          # every weight-gradient block retired exactly once
    assert abs(float(met[0]) / B - loss.item()) < 2e-5 and int(round(float(met[1]))) == want_corr
    assert torch.allclose(d_h.cpu(), h.grad, atol=1e-6, rtol=1e-4)
    assert torch.allclose(d_w.cpu(), fc.weight.grad, atol=1e-6, rtol=1e-4)
    assert torch.allclose(d_b.cpu(), fc.bias.grad, atol=1e-6, rtol=1e-4)


def _ref_step_grads(model, x, y):
    model.zero_grad()
    out = model(x)
    loss = nn.functional.cross_entropy(out, y)
    loss.backward()
    return loss.item(), int((out.argmax(1) == y).sum())


# (B, T, I, H, L): ragged everything; K tails (I = 9 and 20 are not multiples of the 64-wide K block, T*B is not a
# multiple of 8), the scaled geometry's hidden size at T >= 64, and a batch above the persistent kernels' 128 rows
@pytest.mark.parametrize("B,T,I,H,L", [(12, 9, 9, 64, 2), (33, 7, 20, 128, 2), (16, 64, 9, 1024, 2), (150, 5, 9, 64, 1)])
def test_layered_engine_gradients_and_loss_match_torch(B, T, I, H, L):
    torch.manual_seed(H + B)
    model = MotionModel(I, H, L, 6)
    ref = copy.deepcopy(model)
    N = B + 9
    feats, labels = torch.randn(N, T, I), torch.randint(0, 6, (N, 1))
    idx = torch.randperm(N)[:B]
    want_loss, want_corr = _ref_step_grads(ref, feats[idx], labels[idx].view(-1))
    model = model.to(DEV)
    eng = LayeredTrainEngine(model, lr=0.0, timeline=True)      # lr 0: parameters stay put, arena.grad is the raw gradient
    m = eng.train_step_gather(feats.to(DEV), labels.to(DEV), idx.to(DEV)).cpu()
    torch.cuda.synchronize()
    _native.assert_device_healthy()
    assert eng.reducer_status() == 0
    assert abs(float(m[0]) / B - want_loss) < 1e-2 * max(1.0, abs(want_loss)), (float(m[0]) / B, want_loss)
    assert abs(int(round(float(m[1]))) - want_corr) <= max(1, B // 20)
    a = eng.arena
    for (name, p), q in zip(model.named_parameters(), ref.parameters()):
        o, k, shp = a.index[name]
        got = a.grad[o:o + k].view(shp).cpu()
        scale = float(q.grad.abs().max()) + 1e-8
        err = float((got - q.grad).abs().max())
        assert err <= 2e-2 * scale, (name, err, scale)            # bf16 operands, fp32 accumulation
    tl = eng.last_timeline()
    assert tl["status"] == 0
    for name, b in tl["buckets"].items():
        assert b["trigger_us"] is not None and b["collective_done_us"] is not None, (name, b)
        assert b["collective_done_us"] >= b["trigger_us"]
        if name != "fc":   # a layer's bucket can only fire after that layer's backward has been issued and run
            assert b["trigger_us"] >= b["backward_begin_us"]


def test_layered_engine_trains_like_torch_adam():
    torch.manual_seed(3)
    B, T, H, L, steps, lr = 24, 12, 128, 2, 4, 0.0025
    model = MotionModel(9, H, L, 6)
    ref = copy.deepcopy(model)
    xs = [torch.randn(B, T, 9) for _ in range(steps)]
    ys = [torch.randint(0, 6, (B,)) for _ in range(steps)]
    opt = torch.optim.Adam(ref.parameters(), lr=lr)
    want = []
    for s in range(steps):
        opt.zero_grad()
        loss = nn.functional.cross_entropy(ref(xs[s]), ys[s])
        loss.backward()
        opt.step()
        want.append(loss.item())
    model = model.to(DEV)
    eng = LayeredTrainEngine(model, lr=lr)
    got = [float(eng.train_step(xs[s].to(DEV), ys[s].to(DEV)).cpu()[0]) / B for s in range(steps)]
    torch.cuda.synchronize()
    for g, w in zip(got, want):
        assert abs(g - w) < 2e-2 * max(1.0, abs(w)), (got, want)
    # Adam's first steps move every weight by ~lr whatever the gradient's magnitude, so a sign flip of a
    # near-zero bf16 gradient costs up to 2*lr per step on that element: compare in the mean
    diff = torch.cat([(p.detach().cpu() - q.detach()).abs().view(-1) for p, q in zip(model.parameters(), ref.parameters())])
    assert float(diff.mean()) < 0.25 * lr * steps, float(diff.mean())
    sd = eng.optimizer.state_dict()
    assert float(sd["state"][0]["step"]) == steps
    # evaluation path of the same engine
    with torch.no_grad():
        out = ref(xs[0])
        wl = nn.functional.cross_entropy(out, ys[0], reduction="sum").item()
    ev = eng.evaluate(xs[0].to(DEV), ys[0].to(DEV)).cpu()
    assert abs(float(ev[0]) - wl) < 5e-2 * max(1.0, abs(wl))


def test_trainer_picks_layered_engine_for_large_hidden(tmp_path):
    from pytorch_distributed_rnn_b200.data.dataset import MotionDataset
    from pytorch_distributed_rnn_b200.trainer.base import Trainer
    torch.manual_seed(0)
    ds = MotionDataset(torch.randn(96, 16, 9), torch.randint(0, 6, (96, 1)))
    tr = Trainer(MotionModel(9, 64, 2, 6), ds, 48, 0.0025, device=torch.device(DEV))
    assert isinstance(tr.engine, LayeredTrainEngine)
    _, hist, _ = tr.train(2)
    assert len(hist) == 2 and hist[1] < hist[0]
