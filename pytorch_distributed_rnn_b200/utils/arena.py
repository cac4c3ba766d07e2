"""Flat parameter / gradient arenas.

Every parameter of a module becomes a view into ONE contiguous fp32 buffer (and
its ``.grad`` a view into a second one), in ``module.parameters()`` order.  This is
what lets the fused optimizer and the in-kernel all-reduce treat the whole model as
a single vector — the job c10d's Reducer does with flattened buckets for the
reference's ``DistributedDataParallel(model)`` (src/motion/trainer/ddp.py:19) and
Horovod's fusion buffer does for ``hvd.DistributedOptimizer``
(src/motion/trainer/horovod.py:33-35).  ``state_dict()`` is unaffected: names,
shapes and dtypes stay torch's.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch
from torch import nn


class ParamArena:
    """Owns ``flat`` (params) and ``grad`` (grads) and the name -> (offset, numel, shape) map."""

    def __init__(self, module: nn.Module, flat: Optional[torch.Tensor] = None,
                 grad: Optional[torch.Tensor] = None, extra_slots: int = 2):
        params = [(n, p) for n, p in module.named_parameters()]
        if not params:
            raise ValueError("module has no parameters")
        dev = params[0][1].device
        for n, p in params:
            if p.dtype != torch.float32:
                raise TypeError(f"parameter {n} is {p.dtype}; the arena is fp32 (master weights)")
            if p.device != dev:
                raise ValueError("all parameters must live on one device")
        self.module = module
        self.device = dev
        self.index: Dict[str, Tuple[int, int, torch.Size]] = {}
        off = 0
        for n, p in params:
            self.index[n] = (off, p.numel(), p.shape)
            off += p.numel()
        self.numel = off
        self.extra_slots = extra_slots
        # `flat` may be a caller-provided buffer (e.g. symmetric / peer-mapped memory)
        self.flat = flat if flat is not None else torch.empty(off, dtype=torch.float32, device=dev)
        self.grad = grad if grad is not None else torch.zeros(off + extra_slots, dtype=torch.float32, device=dev)
        if self.flat.numel() < off or self.grad.numel() < off:
            raise ValueError("provided arena buffers are too small")
        with torch.no_grad():
            for n, p in params:
                o, k, shp = self.index[n]
                view = self.flat[o:o + k].view(shp)
                view.copy_(p.data)
                p.data = view
                p.grad = self.grad[o:o + k].view(shp)

    # -- helpers --
    def offset(self, name: str) -> int:
        return self.index[name][0]

    def names(self) -> List[str]:
        return list(self.index.keys())

    def zero_grad(self) -> None:
        self.grad.zero_()

    def rebind(self) -> None:
        """Re-point .data/.grad at the arena (after e.g. load_state_dict(assign=True) or .to())."""
        with torch.no_grad():
            for n, p in self.module.named_parameters():
                o, k, shp = self.index[n]
                if p.data.data_ptr() != self.flat[o:o + k].data_ptr():
                    self.flat[o:o + k].view(shp).copy_(p.data)
                    p.data = self.flat[o:o + k].view(shp)
                p.grad = self.grad[o:o + k].view(shp)


def params_checksum(module: nn.Module) -> float:
    """Sum of all parameters — the quantity the reference prints to eyeball rank agreement
    (src/example/example_ddp.py:40-51,83-87)."""
    with torch.no_grad():
        return float(sum(p.double().sum() for p in module.parameters()))
