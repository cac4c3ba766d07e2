"""Large-hidden LSTM / GRU path against an oracle with the SAME operand precision.

`test_rnn_large_gpu.py` compares with fp32 torch and therefore needs a 3-6 % tolerance (bf16 operands), which would
hide a dropped K tail or a short bias sum.  Here the oracle is a plain torch loop in fp32 whose matrix products round
their operands to bf16 exactly where the kernels do (forward: input and hidden state times W; backward: gate gradients
times W / times the activations), so what is left is summation order and a handful of flipped bf16 roundings, and the
tolerances are an order of magnitude tighter.  Also drives the scaled geometry (hidden 1024, batch 64, 2 layers) over a
longer sequence than the fp32 comparison does.
"""
import pytest
import torch

from pytorch_distributed_rnn_b200 import FusedGRU, FusedLSTM
from pytorch_distributed_rnn_b200.ops.gemm import gemm_status, persist_status

pytestmark = pytest.mark.gpu


def _bf(t):
    return t.to(torch.bfloat16).to(torch.float32)


class _Bf16Linear(torch.autograd.Function):
    """y = bf(a) @ bf(w)^T in fp32; gradients from bf(g): da = bf(g) @ bf(w), dw = bf(g)^T @ bf(a)."""

    @staticmethod
    def forward(ctx, a, w):
        ab, wb = _bf(a), _bf(w)
        ctx.save_for_backward(ab, wb)
        return ab @ wb.t()

    @staticmethod
    def backward(ctx, g):
        ab, wb = ctx.saved_tensors
        gb = _bf(g)
        return gb @ wb, gb.reshape(-1, gb.shape[-1]).t() @ ab.reshape(-1, ab.shape[-1])


def _oracle(cell, x, params, L, H):
    """x [B,T,I]; params: list per layer of (w_ih, w_hh, b_ih, b_hh) requiring grad.  Returns out [B,T,H], h_n, c_n."""
    lin = _Bf16Linear.apply
    B, T, _ = x.shape
    inp, hn, cn = x, [], []
    for l in range(L):
        w_ih, w_hh, b_ih, b_hh = params[l]
        gx = lin(inp, w_ih) + b_ih                       # [B,T,G*H]
        h = x.new_zeros(B, H)
        c = x.new_zeros(B, H)
        outs = []
        for t in range(T):
            gh = lin(h, w_hh) + b_hh
            if cell == "lstm":
                i, f, g, o = (gx[:, t] + gh).chunk(4, 1)
                c = torch.sigmoid(f) * c + torch.sigmoid(i) * torch.tanh(g)
                h = torch.sigmoid(o) * torch.tanh(c)
            else:
                xr, xz, xn = gx[:, t].chunk(3, 1)
                hr, hz, hnn = gh.chunk(3, 1)
                r, z = torch.sigmoid(xr + hr), torch.sigmoid(xz + hz)
                n = torch.tanh(xn + r * hnn)
                h = (1 - z) * n + z * h
            outs.append(h)
        inp = torch.stack(outs, 1)
        hn.append(h)
        cn.append(c)
    return inp, torch.stack(hn), torch.stack(cn)


def _rel(a, b):
    return float((a - b).abs().max()) / (float(b.abs().max()) + 1e-12)


@pytest.mark.parametrize("cell,I,H,L,B,T", [
    ("lstm", 9, 64, 2, 23, 11),       # T*B = 253 rows: not a multiple of any tile or bias-sum slab
    ("lstm", 40, 128, 1, 80, 9),      # M = 128 persistent variants
    ("lstm", 9, 1024, 2, 64, 48),     # the scaled geometry, 2 layers, 48 steps
    ("lstm", 20, 192, 1, 130, 5),     # per-timestep formulation (batch > 128)
    ("gru", 9, 64, 2, 23, 11),
    ("gru", 9, 1024, 1, 64, 24),      # persistent GRU at the scaled geometry
    ("gru", 24, 256, 2, 80, 9),       # 2-D backward, M = 128
])
def test_large_path_matches_bf16_operand_oracle(cell, I, H, L, B, T):
    torch.manual_seed(1000 + H + T)
    cls = FusedLSTM if cell == "lstm" else FusedGRU
    m = cls(I, H, L, batch_first=True).cuda()
    x = torch.randn(B, T, I, device="cuda")
    wo = torch.randn(B, T, H, device="cuda") / T ** 0.5
    xd = x.clone().requires_grad_()
    out, st = m(xd)
    h = st[0] if isinstance(st, tuple) else st
    loss = (out * wo).sum() + h.sum()
    if isinstance(st, tuple):
        loss = loss + 0.5 * st[1].sum()
    loss.backward()

    ref_params = [[getattr(m, f"{n}_l{l}").detach().clone().requires_grad_() for n in ("weight_ih", "weight_hh", "bias_ih", "bias_hh")]
                  for l in range(L)]
    xr = x.clone().requires_grad_()
    ro, rh, rc = _oracle(cell, xr, ref_params, L, H)
    rloss = (ro * wo).sum() + rh.sum()
    if isinstance(st, tuple):
        rloss = rloss + 0.5 * rc.sum()
    rloss.backward()
    torch.cuda.synchronize()
    assert gemm_status() == 0 and persist_status() == 0

    errs = {"out": _rel(out.detach(), ro.detach()), "h_n": _rel(h.detach(), rh.detach()), "dx": _rel(xd.grad, xr.grad)}
    if isinstance(st, tuple):
        errs["c_n"] = _rel(st[1].detach(), rc.detach())
    for l in range(L):
        for n, q in zip(("weight_ih", "weight_hh", "bias_ih", "bias_hh"), ref_params[l]):
            errs[f"{n}_l{l}"] = _rel(getattr(m, f"{n}_l{l}").grad, q.grad)
    print(f"bf16-oracle {cell} I={I} H={H} L={L} B={B} T={T}: " + " ".join(f"{k}={v:.2e}" for k, v in errs.items()))
    # measured on B200 (profiles/r2_final/pytest_oracle_loss.log): states <= 3.4e-4, dx / weights <= 1.1e-3, biases <= 2.6e-5
    for k, v in errs.items():
        tol = 1e-3 if k in ("out", "h_n", "c_n") else (2e-4 if k.startswith("bias") else 3e-3)
        assert v < tol, (k, v, errs)
