"""Shared pieces of the toy plumbing examples (E1-E4): the 24-sample regression task and one pass over it
(the reference repeats them in each script: src/example/example_ddp.py:11-32,40-92, example_horovod.py,
example_single.py).

Each example only differs in HOW gradients are synchronised (not at all / DDP wrapper / Horovod-style optimizer);
the optimisation pass and the lines it prints (``parameters:`` is what the tests compare across ranks) live here.
"""
from __future__ import annotations

from typing import Callable, Optional

import torch
from torch import nn
from torch.utils.data import DataLoader

import _path  # noqa: F401
from pytorch_distributed_rnn_b200 import ToyDataset, params_checksum

GLOBAL_BATCH = 12
LR = 0.001


def toy_loader(world_size: int, sampler=None, size: int = 24) -> DataLoader:
    return DataLoader(dataset=ToyDataset(size=size), batch_size=GLOBAL_BATCH // world_size, sampler=sampler)


def reporter(rank: int, quiet: bool = False) -> Callable[..., None]:
    if quiet:
        return lambda *a: None
    return lambda tag, value: print("rank", rank, f"{tag}:", value)


def training_pass(net: nn.Module, optimizer, loader, device, say: Callable[..., None],
                  verbose_batches: bool = False, grad_of: Optional[torch.Tensor] = None) -> None:
    """One epoch of MSE regression; prints the parameter checksum after every step."""
    criterion = nn.MSELoss()
    for inputs, labels in loader:
        inputs, labels = inputs.to(device), labels.to(device)
        optimizer.zero_grad()
        prediction = net(inputs)
        if verbose_batches:
            say("inputs", float(inputs.sum()))
            say("labels", float(labels.sum()))
        else:
            say("batch", float(inputs.sum() + labels.sum()))
        loss = criterion(prediction, labels)
        say("loss", loss.item())
        loss.backward()
        optimizer.step()
        say("parameters", params_checksum(net))
        if grad_of is not None:
            say("grad", float(grad_of.grad.sum()) if grad_of.grad is not None else None)
