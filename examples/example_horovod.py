#!/usr/bin/env python
"""E4 — Horovod-style toy (parity: src/example/example_horovod.py): broadcast_parameters +
DistributedOptimizer(SGD) + a sharded sampler.
Launch:  python -m pytorch_distributed_rnn_b200.parallel.launch -n 2 examples/example_horovod.py
"""
import numpy as np
import torch
from torch import nn, optim
from torch.utils.data import DataLoader

import _path  # noqa: F401
from pytorch_distributed_rnn_b200 import ToyDataset, ToyModel, params_checksum
from pytorch_distributed_rnn_b200.data.sampler import ShardedEpochSampler
from pytorch_distributed_rnn_b200.parallel import horovod as hvd
from pytorch_distributed_rnn_b200.parallel import process_group as pg


def run(rank, world_size, device, quiet=False):
    say = (lambda *a: None) if quiet else print
    model = ToyModel().to(device)
    say("rank ", rank, "initial:", params_checksum(model))
    hvd.broadcast_parameters(model.state_dict(), root_rank=0)
    say("rank", rank, "synced:", params_checksum(model))
    loss_fn = nn.MSELoss()
    optimizer = hvd.DistributedOptimizer(optim.SGD(model.parameters(), lr=0.001),
                                         named_parameters=model.named_parameters())
    dataset = ToyDataset(size=24)
    batch_size = 12 // world_size
    loader = DataLoader(dataset=dataset, batch_size=batch_size,
                        sampler=ShardedEpochSampler(len(dataset), world_size, rank))
    for inputs, labels in loader:
        inputs, labels = inputs.to(device), labels.to(device)
        optimizer.zero_grad()
        say("rank", rank, "inputs:", float(inputs.sum()))
        outputs = model(inputs)
        say("rank", rank, "labels:", float(labels.sum()))
        loss_fn(outputs, labels).backward()
        optimizer.step()
        say("rank", rank, "parameters:", params_checksum(model))
    return model


if __name__ == "__main__":
    np.random.seed(0)
    torch.manual_seed(0)
    hvd.init()
    run(hvd.rank(), hvd.size(), pg.rank_info().device)
