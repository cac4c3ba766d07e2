"""Large-hidden LSTM / GRU path (tcgen05 GEMMs + fused cell kernels, bf16 operands) vs fp32 torch."""
import pytest
import torch
from torch import nn

from pytorch_distributed_rnn_b200 import FusedGRU, FusedLSTM
from pytorch_distributed_rnn_b200.ops.gemm import gemm_status, persist_status

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return float((a - b).abs().max()) / (float(b.abs().max()) + 1e-6)


@pytest.mark.parametrize("cls,stock,I,H,L,B,T", [
    (FusedLSTM, nn.LSTM, 9, 64, 2, 24, 12),
    (FusedLSTM, nn.LSTM, 40, 128, 1, 16, 20),
    (FusedLSTM, nn.LSTM, 64, 256, 2, 8, 6),
    (FusedLSTM, nn.LSTM, 24, 128, 2, 80, 9),    # 64 < B <= 128: the M = 128 persistent variants
    (FusedLSTM, nn.LSTM, 16, 1024, 1, 64, 5),   # the scaled configuration's geometry (128 CTAs, 64 KB W shards)
    (FusedLSTM, nn.LSTM, 16, 192, 1, 130, 4),   # B > 128: per-timestep formulation
    (FusedGRU, nn.GRU, 9, 64, 2, 24, 12),
    (FusedGRU, nn.GRU, 9, 40, 2, 16, 10),      # H % 8 == 0 but outside the fused small-hidden family
    (FusedGRU, nn.GRU, 48, 128, 1, 8, 7),
    (FusedGRU, nn.GRU, 24, 256, 2, 80, 9),     # persistent GRU, 2-D backward (3H % 256 == 0), M = 128 variants
    (FusedGRU, nn.GRU, 16, 1024, 1, 64, 5),    # persistent GRU at the scaled geometry
    (FusedGRU, nn.GRU, 16, 192, 1, 130, 4),    # B > 128: per-timestep formulation
])
def test_large_path_forward_backward(cls, stock, I, H, L, B, T):
    torch.manual_seed(H + T)
    m = cls(I, H, L, batch_first=True).cuda()
    ref = stock(I, H, L, batch_first=True)
    ref.load_state_dict({k: v.detach().cpu() for k, v in m.state_dict().items()})
    x = torch.randn(B, T, I)
    wo = torch.randn(B, T, H)
    xd = x.cuda().requires_grad_()
    out, st = m(xd)
    h = st[0] if isinstance(st, tuple) else st
    ((out * wo.cuda()).sum() + h.sum()).backward()
    xr = x.clone().requires_grad_()
    ro, rst = ref(xr)
    rh = rst[0] if isinstance(rst, tuple) else rst
    ((ro * wo).sum() + rh.sum()).backward()
    torch.cuda.synchronize()
    assert gemm_status() == 0
    assert persist_status() == 0
    assert _rel(out.detach().cpu(), ro.detach()) < 3e-2
    assert _rel(h.detach().cpu(), rh.detach()) < 3e-2
    if isinstance(st, tuple):
        assert _rel(st[1].detach().cpu(), rst[1].detach()) < 3e-2
    assert _rel(xd.grad.cpu(), xr.grad) < 6e-2
    for (n, p), q in zip(m.named_parameters(), ref.parameters()):
        assert _rel(p.grad.cpu(), q.grad) < 6e-2, n
