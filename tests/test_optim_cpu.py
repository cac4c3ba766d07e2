"""Flat optimizers vs torch.optim on CPU + stock state_dict interchange."""
import copy

import torch
from torch import nn

from pytorch_distributed_rnn_b200 import FusedAdam, FusedSGD, MotionModel, ParamArena


def _train(model, opt, steps=5):
    torch.manual_seed(1)
    for _ in range(steps):
        x, y = torch.randn(8, 6, 9), torch.randint(0, 6, (8,))
        opt.zero_grad()
        nn.functional.cross_entropy(model(x), y).backward()
        opt.step()


def test_adam_matches_torch():
    torch.manual_seed(0)
    a = MotionModel(9, 32, 2, 6)
    b = copy.deepcopy(a)
    _train(a, FusedAdam(ParamArena(a), lr=0.0025))
    _train(b, torch.optim.Adam(b.parameters(), lr=0.0025))
    for (n, p), q in zip(a.named_parameters(), b.parameters()):
        assert torch.allclose(p, q, atol=2e-6), n


def test_sgd_momentum_matches_torch():
    torch.manual_seed(0)
    a = MotionModel(9, 32, 1, 6)
    b = copy.deepcopy(a)
    _train(a, FusedSGD(ParamArena(a), lr=0.01, momentum=0.9))
    _train(b, torch.optim.SGD(b.parameters(), lr=0.01, momentum=0.9))
    for p, q in zip(a.parameters(), b.parameters()):
        assert torch.allclose(p, q, atol=2e-6)


def test_state_dict_interchange_with_stock_adam():
    torch.manual_seed(0)
    a = MotionModel(9, 32, 2, 6)
    fa = FusedAdam(ParamArena(a), lr=0.0025)
    _train(a, fa, 3)
    sd = fa.state_dict()
    b = copy.deepcopy(a)
    stock = torch.optim.Adam(b.parameters(), lr=1.0)
    stock.load_state_dict(sd)  # ours -> stock
    assert stock.param_groups[0]["lr"] == 0.0025
    _train(a, fa, 2)
    _train(b, stock, 2)
    for p, q in zip(a.parameters(), b.parameters()):
        assert torch.allclose(p, q, atol=2e-6)
    c = copy.deepcopy(b)
    fc = FusedAdam(ParamArena(c), lr=1.0)
    fc.load_state_dict(stock.state_dict())  # stock -> ours
    assert fc.step_count == 5 and fc.param_groups[0]["lr"] == 0.0025
    _train(b, stock, 1)
    _train(c, fc, 1)
    for p, q in zip(b.parameters(), c.parameters()):
        assert torch.allclose(p, q, atol=2e-6)
