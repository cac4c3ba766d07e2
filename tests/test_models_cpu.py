"""Model parity on CPU: init stream, state_dict contract, composite oracle vs ATen."""
import pytest
import torch
from torch import nn

from pytorch_distributed_rnn_b200 import FusedGRU, FusedLSTM, MotionModel, ParamArena
from pytorch_distributed_rnn_b200.models.rnn import rnn_composite


class RefMotionModel(nn.Module):  # stock-torch twin of src/motion/model.py
    def __init__(self, i, h, l, o):
        super().__init__()
        self.lstm = nn.LSTM(i, h, l, batch_first=True)
        self.fc = nn.Linear(h, o)

    def forward(self, x):
        out, _ = self.lstm(x)
        return self.fc(out[:, -1, :])


def test_same_seed_same_init_and_keys():
    torch.manual_seed(123456789)
    ref = RefMotionModel(9, 32, 2, 6)
    torch.manual_seed(123456789)
    ours = MotionModel(9, 32, 2, 6)
    rs, os_ = ref.state_dict(), ours.state_dict()
    assert list(rs.keys()) == list(os_.keys())
    for k in rs:
        assert rs[k].shape == os_[k].shape and rs[k].dtype == os_[k].dtype
        assert torch.equal(rs[k], os_[k]), k
    assert sum(p.numel() for p in ours.parameters()) == 14150


def test_state_dict_roundtrip_into_stock_modules():
    ours = MotionModel(9, 32, 2, 6)
    ParamArena(ours)  # arena views must not change what state_dict exposes
    ref = RefMotionModel(9, 32, 2, 6)
    ref.load_state_dict(ours.state_dict())
    x = torch.randn(5, 16, 9)
    assert torch.allclose(ref(x), ours(x), atol=1e-6)
    ours2 = MotionModel(9, 32, 2, 6)
    ours2.load_state_dict(ref.state_dict())
    assert torch.allclose(ref(x), ours2(x), atol=1e-6)


@pytest.mark.parametrize("cls,cell,stock", [(FusedLSTM, "lstm", nn.LSTM), (FusedGRU, "gru", nn.GRU)])
def test_composite_matches_aten(cls, cell, stock):
    torch.manual_seed(0)
    m = cls(7, 12, 3, batch_first=True)
    s = stock(7, 12, 3, batch_first=True)
    s.load_state_dict(m.state_dict())
    x = torch.randn(4, 9, 7)
    out_a, st_a = m(x)
    out_s, st_s = s(x)
    out_c, st_c = rnn_composite(cell, x, m.flat_weights(), 3, 12)
    assert torch.allclose(out_a, out_s, atol=1e-6)
    assert torch.allclose(out_c, out_s, atol=1e-5)
    if cell == "lstm":
        assert torch.allclose(st_c[0], st_s[0], atol=1e-5) and torch.allclose(st_c[1], st_s[1], atol=1e-5)
    else:
        assert torch.allclose(st_c, st_s, atol=1e-5)


def test_arena_views_share_memory():
    m = MotionModel(9, 32, 2, 6)
    a = ParamArena(m)
    assert a.numel == 14150 and a.grad.numel() == 14152
    a.flat.zero_()
    assert all(float(p.detach().abs().sum()) == 0 for p in m.parameters())
    m(torch.randn(2, 4, 9)).sum().backward()
    assert float(a.grad[:a.numel].abs().sum()) > 0  # autograd accumulated into the arena views


def test_rnn_modules_expose_the_stock_attribute_surface():
    """Code written against nn.LSTM / nn.GRU (flatten_parameters, all_weights, bias, ...) runs unchanged."""
    for cls, stock in ((FusedLSTM, nn.LSTM), (FusedGRU, nn.GRU)):
        m, ref = cls(9, 32, 2, batch_first=True), stock(9, 32, 2, batch_first=True)
        m.flatten_parameters()
        assert [[tuple(t.shape) for t in layer] for layer in m.all_weights] == \
               [[tuple(t.shape) for t in layer] for layer in ref.all_weights]
        for attr in ("input_size", "hidden_size", "num_layers", "bias", "batch_first", "dropout", "bidirectional", "proj_size"):
            assert getattr(m, attr) == getattr(ref, attr), attr
