#!/usr/bin/env python
"""E4 — the Horovod-flavoured toy (capability parity: src/example/example_horovod.py): parameters are
broadcast from rank 0, the optimizer averages gradients at ``step()``, every rank reads its own shard.
Launch:  python -m pytorch_distributed_rnn_b200.parallel.launch -n 2 examples/example_horovod.py
"""
import numpy as np
import torch
from torch import optim

import _toy
from pytorch_distributed_rnn_b200 import ToyModel, params_checksum
from pytorch_distributed_rnn_b200.data.sampler import ShardedEpochSampler
from pytorch_distributed_rnn_b200.parallel import horovod as hvd
from pytorch_distributed_rnn_b200.parallel import process_group as pg


def run(rank, world_size, device, quiet=False):
    say = _toy.reporter(rank, quiet)
    net = ToyModel().to(device)
    say("initial", params_checksum(net))
    hvd.broadcast_parameters(net.state_dict(), root_rank=0)
    say("synced", params_checksum(net))
    averaging_sgd = hvd.DistributedOptimizer(optim.SGD(net.parameters(), lr=_toy.LR),
                                             named_parameters=net.named_parameters())
    loader = _toy.toy_loader(world_size, sampler=ShardedEpochSampler(24, world_size, rank))
    _toy.training_pass(net, averaging_sgd, loader, device, say, verbose_batches=True)
    return net


if __name__ == "__main__":
    np.random.seed(0)
    torch.manual_seed(0)
    hvd.init()
    run(hvd.rank(), hvd.size(), pg.rank_info().device)
