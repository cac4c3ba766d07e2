#!/usr/bin/env python
"""E1 — the smallest plumbing check: one fused-SGD step of ``Linear(10, 10)`` in a single process
(capability parity: src/example/example_single.py), on the GPU when there is one."""
import torch
from torch import nn

import _path  # noqa: F401
from pytorch_distributed_rnn_b200 import FusedSGD, ParamArena


def one_step(net: nn.Module, optimizer, device) -> float:
    x, target = torch.randn(20, 10, device=device), torch.randn(20, 10, device=device)
    loss = nn.functional.mse_loss(net(x), target)
    loss.backward()
    optimizer.step()
    return float(loss.detach())


def run(device="cpu"):
    net = nn.Linear(10, 10).to(device)
    one_step(net, FusedSGD(ParamArena(net), lr=0.001), device)
    print(list(net.parameters()))
    return net


if __name__ == "__main__":
    run("cuda" if torch.cuda.is_available() else "cpu")
