#!/usr/bin/env python
"""E1 — one SGD step of Linear(10,10), single process (parity: src/example/example_single.py)."""
import torch
from torch import nn

import _path  # noqa: F401
from pytorch_distributed_rnn_b200 import FusedSGD, ParamArena


def run(device="cpu"):
    model = nn.Linear(10, 10).to(device)
    loss_fn = nn.MSELoss()
    optimizer = FusedSGD(ParamArena(model), lr=0.001)
    outputs = model(torch.randn(20, 10, device=device))
    labels = torch.randn(20, 10, device=device)
    loss_fn(outputs, labels).backward()
    optimizer.step()
    print(list(model.parameters()))
    return model


if __name__ == "__main__":
    run("cuda" if torch.cuda.is_available() else "cpu")
