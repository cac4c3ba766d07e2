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
from torch import optim

import _toy
from pytorch_distributed_rnn_b200 import ToyModel, params_checksum
from pytorch_distributed_rnn_b200.data.sampler import ShardedEpochSampler
from pytorch_distributed_rnn_b200.parallel import process_group as pg
from pytorch_distributed_rnn_b200.parallel.ddp import DistributedDataParallel as DDP


def run(rank, world_size, device, sharded=False, quiet=False):
    say = _toy.reporter(rank, quiet)
    torch.manual_seed(0)
    np.random.seed(0)
    net = ToyModel().to(device)
    say("initial", params_checksum(net))
    replica = DDP(net)     # broadcasts rank 0's parameters; averages gradients bucket by bucket during backward
    say("synced", params_checksum(replica))
    sgd = optim.SGD(replica.parameters(), lr=_toy.LR)
    shard = ShardedEpochSampler(24, world_size, rank, shuffle=False) if sharded else None
    _toy.training_pass(replica, sgd, _toy.toy_loader(world_size, sampler=shard), device, say,
                       grad_of=sgd.param_groups[0]["params"][0])
    return replica


if __name__ == "__main__":
    torch.manual_seed(0)
    np.random.seed(0)
    info = pg.init_distributed()
    run(info.rank, info.world_size, info.device, sharded="--sharded" in sys.argv)
