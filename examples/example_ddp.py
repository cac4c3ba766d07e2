#!/usr/bin/env python
"""E3 — the DDP toy entry point (parity: src/example/example_ddp.py).

ToyModel = Linear(10,10)->ReLU->Linear(10,5), ToyDataset(24), MSELoss, SGD(lr=1e-3),
batch = 12 // world_size, NO sampler (every rank iterates the same 24 samples — quirk Q18,
kept; pass --sharded to also exercise a sharded sampler).  Pass criterion (README.md:9): the
last ``parameters:`` value is identical on every rank.
Launch:  python -m pytorch_distributed_rnn_b200.parallel.launch -n 2 examples/example_ddp.py
"""
import sys

import numpy as np
import torch
from torch import nn, optim
from torch.utils.data import DataLoader

import _path  # noqa: F401
from pytorch_distributed_rnn_b200 import ToyDataset, ToyModel, params_checksum
from pytorch_distributed_rnn_b200.data.sampler import ShardedEpochSampler
from pytorch_distributed_rnn_b200.parallel import process_group as pg
from pytorch_distributed_rnn_b200.parallel.ddp import DistributedDataParallel as DDP


def run(rank, world_size, device, sharded=False, quiet=False):
    say = (lambda *a: None) if quiet else print
    torch.manual_seed(0)
    np.random.seed(0)
    model = ToyModel().to(device)
    say("rank", rank, "initial:", params_checksum(model))
    ddp_model = DDP(model)
    say("rank", rank, "synced:", params_checksum(ddp_model))
    loss_fn = nn.MSELoss()
    optimizer = optim.SGD(ddp_model.parameters(), lr=0.001)
    optimizer.zero_grad()
    dataset = ToyDataset(size=24)
    batch_size = 12 // world_size
    sampler = ShardedEpochSampler(len(dataset), world_size, rank, shuffle=False) if sharded else None
    loader = DataLoader(dataset=dataset, batch_size=batch_size, sampler=sampler)
    for inputs, labels in loader:
        inputs, labels = inputs.to(device), labels.to(device)
        optimizer.zero_grad()
        outputs = ddp_model(inputs)
        say("rank", rank, "batch:", float(inputs.sum() + labels.sum()))
        loss = loss_fn(outputs, labels)
        say("rank", rank, "loss:", loss.item())
        loss.backward()
        optimizer.step()
        say("rank", rank, "parameters:", params_checksum(ddp_model))
        g = optimizer.param_groups[0]["params"][0].grad
        say("rank", rank, "grad:", float(g.sum()) if g is not None else None)
    return ddp_model


if __name__ == "__main__":
    torch.manual_seed(0)
    np.random.seed(0)
    info = pg.init_distributed()
    run(info.rank, info.world_size, info.device, sharded="--sharded" in sys.argv)
