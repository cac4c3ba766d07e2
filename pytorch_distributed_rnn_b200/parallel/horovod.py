"""Horovod-style synchronous data parallelism: the all-reduce lives in the optimizer.

API surface the reference uses from ``horovod.torch`` (src/motion/trainer/horovod.py:17-42,
src/example/example_horovod.py:42-53,80-83): ``init``, ``rank``, ``size``, ``local_rank``,
``broadcast_parameters(state_dict, root_rank)``, ``DistributedOptimizer(optimizer,
named_parameters)``, plus the rest of the everyday ``horovod.torch`` vocabulary so that other scripts written
against it keep working: ``allreduce`` / ``allreduce_`` (``op=Average | Sum``), ``broadcast`` / ``broadcast_``,
``allgather``, ``broadcast_optimizer_state``, ``local_size``, ``is_initialized``, ``barrier`` / ``join``, ``shutdown``.

Horovod's C++ core negotiates tensor readiness by name through a rank-0 coordinator and
packs ready tensors into a fusion buffer each cycle.  Training graphs here are static, so
no negotiation is needed: the fusion plan is fixed at construction (all named gradients,
one buffer in parameter order), gradient hooks fire the fused all-reduce (op = Average) as
soon as the last gradient of the step is produced, and ``step()`` is the synchronisation
point — same semantics, zero coordinator round-trips.  The model is NOT wrapped, so
``state_dict()`` keys stay un-prefixed (contrast :mod:`.ddp`).
"""
from __future__ import annotations

from typing import Iterable, Optional, Tuple

import torch
import torch.distributed as dist
from torch._utils import _flatten_dense_tensors, _unflatten_dense_tensors

from . import process_group as pg
from .comm import TorchDistComm, make_comm

_comm = None


def init(comm=None):
    global _comm
    pg.init_distributed()
    _comm = comm if comm is not None else make_comm()
    return _comm


def _c():
    if _comm is None:
        init()
    return _comm


def rank() -> int:
    return _c().rank


def size() -> int:
    return _c().world


def local_rank() -> int:
    return pg.rank_info().local_rank


def local_size() -> int:
    """Ranks on this node (one per GPU): what the launcher / torchrun export as LOCAL_WORLD_SIZE, else the world."""
    import os
    return int(os.environ.get("LOCAL_WORLD_SIZE", os.environ.get("OMPI_COMM_WORLD_LOCAL_SIZE", size())))


def is_initialized() -> bool:
    return _comm is not None


def shutdown() -> None:
    """Release the communicator (symmetric memory, peer mappings) and the process group."""
    global _comm
    if _comm is not None and hasattr(_comm, "close"):
        _comm.close()
    _comm = None
    pg.shutdown()


Average, Sum = "average", "sum"     # reduction ops of allreduce (horovod.torch.Average / Sum)


def _want_average(average, op) -> bool:
    if op is not None:
        if op not in (Average, Sum):
            raise ValueError(f"unsupported reduction op {op!r} (Average or Sum)")
        if average is not None:
            raise ValueError("pass either average= or op=, not both")
        return op == Average
    return True if average is None else bool(average)


def allreduce_(tensor: torch.Tensor, average: Optional[bool] = None, op: Optional[str] = None) -> torch.Tensor:
    """In-place all-reduce of a contiguous tensor (default: average)."""
    if not tensor.is_contiguous():
        raise ValueError("allreduce_: tensor must be contiguous")
    _c().allreduce_flat(tensor.detach().view(-1), average=_want_average(average, op))
    return tensor


def allreduce(tensor: torch.Tensor, average: Optional[bool] = None, op: Optional[str] = None) -> torch.Tensor:
    return allreduce_(tensor.detach().clone().contiguous(), average=average, op=op)


def broadcast_(tensor: torch.Tensor, root_rank: int = 0) -> torch.Tensor:
    if size() > 1:
        flat = tensor.detach().contiguous().view(-1)
        _c().broadcast_flat(flat, root=root_rank)
        if flat.data_ptr() != tensor.data_ptr():
            tensor.detach().copy_(flat.view_as(tensor))
    return tensor


def broadcast(tensor: torch.Tensor, root_rank: int = 0) -> torch.Tensor:
    return broadcast_(tensor.detach().clone(), root_rank)


def allgather(tensor: torch.Tensor) -> torch.Tensor:
    """Concatenation of every rank's tensor along dimension 0 (first dimensions may differ, as in Horovod).  A
    convenience outside the training step: it goes through the process group's all_gather."""
    if size() == 1:
        return tensor.detach().clone()
    t = tensor.detach().contiguous()
    n = torch.tensor([t.shape[0] if t.dim() else 1], dtype=torch.int64, device=t.device)
    counts = [torch.zeros_like(n) for _ in range(size())]
    dist.all_gather(counts, n)
    counts = [int(c) for c in counts]
    rows = t.reshape(counts[rank()], -1) if t.dim() else t.reshape(1, 1)
    pad = torch.zeros(max(counts), rows.shape[1], dtype=t.dtype, device=t.device)
    pad[:rows.shape[0]] = rows
    parts = [torch.empty_like(pad) for _ in range(size())]
    dist.all_gather(parts, pad)
    out = torch.cat([p[:c] for p, c in zip(parts, counts)], 0)
    return out.reshape(sum(counts), *t.shape[1:]) if t.dim() else out.reshape(-1)


def barrier() -> None:
    _c().barrier()


join = barrier   # horovod.torch.join(): wait for every rank (there is no uneven-data handling to flush here)


def broadcast_parameters(params, root_rank: int = 0) -> None:
    """Broadcast a ``state_dict()`` or an iterable of ``(name, tensor)`` from ``root_rank`` in place.

    One fused message (S10 in SURVEY §2.5 issues one broadcast per entry; here entries of the
    same dtype are packed)."""
    if isinstance(params, dict):
        items = sorted(params.items())
    else:
        items = sorted(list(params))
    tensors = [t for _, t in items if isinstance(t, torch.Tensor)]
    if not tensors or size() == 1:
        return
    by_dtype = {}
    for t in tensors:
        by_dtype.setdefault((t.dtype, t.device), []).append(t)
    for group in by_dtype.values():
        flat = _flatten_dense_tensors([t.detach() for t in group])
        _c().broadcast_flat(flat, root=root_rank)
        for t, src in zip(group, _unflatten_dense_tensors(flat, group)):
            t.detach().copy_(src)


def broadcast_optimizer_state(optimizer: torch.optim.Optimizer, root_rank: int = 0) -> None:
    tensors = []
    for st in optimizer.state.values():
        for k in sorted(st):
            if isinstance(st[k], torch.Tensor):
                tensors.append((k, st[k]))
    for name in ("exp_avg", "exp_avg_sq"):  # flat optimizers keep state outside .state
        t = getattr(optimizer, name, None)
        if isinstance(t, torch.Tensor):
            tensors.append((name, t))
    broadcast_parameters([(f"{i}.{k}", t) for i, (k, t) in enumerate(tensors)], root_rank)


class DistributedOptimizer:
    """Wraps any ``torch.optim.Optimizer``; averages gradients across ranks before ``step()``.

    ``backward_passes_per_step`` (Horovod's own knob for gradient accumulation): the fused all-reduce is launched by
    the gradient hooks of the LAST of that many backward passes only; with the default of 1 a second ``backward()``
    before ``step()`` is an error instead of silently re-reducing an already reduced buffer."""

    def __init__(self, optimizer: torch.optim.Optimizer, named_parameters: Optional[Iterable[Tuple[str, torch.Tensor]]] = None,
                 backward_passes_per_step: int = 1):
        self._opt = optimizer
        self._comm = _c()
        self._passes_per_step = max(1, int(backward_passes_per_step))
        self._passes = 0
        if named_parameters is not None:
            named = list(named_parameters)
        else:
            named = [(f"p{i}", p) for g in optimizer.param_groups for i, p in enumerate(g["params"])]
        names = [n for n, _ in named]
        if len(set(names)) != len(names):
            raise ValueError("named_parameters must have unique names")
        self._params = [p for _, p in named if p.requires_grad]
        self._ready = 0
        self._work = None
        self._flat = None
        self._synced = True
        if self._comm.world > 1:
            for p in self._params:
                p.register_post_accumulate_grad_hook(self._hook)

    # delegate everything else (param_groups, state_dict, zero_grad, ...) to the wrapped optimizer
    def __getattr__(self, name):
        return getattr(self._opt, name)

    def _hook(self, p):
        if not self._synced:
            raise RuntimeError("backward() ran again after the gradients had been all-reduced and before step(): "
                               "pass backward_passes_per_step=N to DistributedOptimizer to accumulate over N passes")
        self._ready += 1
        if self._ready == len(self._params):
            self._ready = 0
            self._passes += 1
            if self._passes == self._passes_per_step:
                self._launch()

    def _launch(self):
        grads = [p.grad for p in self._params if p.grad is not None]
        arena = getattr(self._opt, "arena", None)
        if arena is not None and all(g.data_ptr() == arena.grad[o:o + 1].data_ptr()
                                     for g, (o, k, s) in zip(grads, arena.index.values())) and len(grads) == len(arena.index):
            self._flat, self._grads = arena.grad[:arena.numel], None  # already one fusion buffer
        else:
            self._flat, self._grads = _flatten_dense_tensors(grads), grads
        if isinstance(self._comm, TorchDistComm):
            self._work = dist.all_reduce(self._flat, async_op=True)
        else:
            self._comm.allreduce_flat(self._flat, average=False)
            self._work = None
        self._synced = False

    def synchronize(self):
        if self._comm.world == 1:
            return
        if self._synced:
            self._launch()  # some gradients never fired (unused parameters / manual grads / fewer passes than announced)
        if not self._synced:
            if self._work is not None:
                self._work.wait()
            self._flat.div_(self._comm.world)
            if self._grads is not None:
                for g, src in zip(self._grads, _unflatten_dense_tensors(self._flat, self._grads)):
                    g.copy_(src)
            self._synced = True
        self._ready = 0
        self._passes = 0

    def step(self, closure=None):
        self.synchronize()
        return self._opt.step(closure)

    def zero_grad(self, *a, **kw):
        return self._opt.zero_grad(*a, **kw)
