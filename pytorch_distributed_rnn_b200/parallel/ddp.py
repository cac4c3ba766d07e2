"""DistributedDataParallel — synchronous data parallelism, reducer-bucket flavour.

Same observable behaviour as the ``DistributedDataParallel(model)`` the reference wraps its
models in (src/motion/trainer/ddp.py:19, src/example/example_ddp.py:46):

* constructing the wrapper broadcasts rank 0's parameters (and buffers) to every rank;
* ``state_dict()`` keys carry the ``module.`` prefix;
* during ``loss.backward()`` gradients are averaged across ranks, bucket by bucket, as soon
  as every gradient of a bucket has been produced (buckets are filled in reverse parameter
  order: 1 MiB first bucket, 25 MiB cap — c10d's defaults), so communication overlaps the
  rest of backward; when ``backward()`` returns every rank holds identical gradients.

Implementation: parameters and gradients live in one flat arena; a bucket is a slice of it,
so "flatten -> all-reduce -> unflatten" is just an all-reduce of a view.  The hooks are
``register_post_accumulate_grad_hook`` + one autograd-engine callback per backward.  On the
fused engine path (MotionModel on CUDA) the trainer bypasses autograd entirely and the
all-reduce is the in-kernel NVLink one (see :mod:`.symm`); this wrapper then only provides
the broadcast, the prefix and the arena.
"""
from __future__ import annotations

from typing import List, Optional

import torch
from torch import nn
from torch.autograd import Variable

from ..utils.arena import ParamArena
from .comm import TorchDistComm, make_comm

FIRST_BUCKET_BYTES = 1 << 20
BUCKET_CAP_BYTES = 25 << 20


def plan_buckets(sizes_bytes: List[int], first_cap: int = FIRST_BUCKET_BYTES, cap: int = BUCKET_CAP_BYTES) -> List[List[int]]:
    """Assign parameter indices to buckets in REVERSE order (gradients become ready back to front)."""
    buckets: List[List[int]] = []
    cur: List[int] = []
    cur_bytes = 0
    limit = first_cap
    for i in reversed(range(len(sizes_bytes))):
        if cur and cur_bytes + sizes_bytes[i] > limit:
            buckets.append(cur)
            cur, cur_bytes, limit = [], 0, cap
        cur.append(i)
        cur_bytes += sizes_bytes[i]
    if cur:
        buckets.append(cur)
    return buckets


class DistributedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, comm=None, arena: Optional[ParamArena] = None,
                 first_bucket_bytes: int = FIRST_BUCKET_BYTES, bucket_cap_bytes: int = BUCKET_CAP_BYTES):
        super().__init__()
        self.module = module
        self.comm = comm if comm is not None else make_comm()
        self.arena = arena if arena is not None else ParamArena(module)
        self.require_backward_grad_sync = True
        # (S8) constructor broadcast: one coalesced message for all parameters, then buffers
        self.comm.broadcast_flat(self.arena.flat[:self.arena.numel], root=0)
        for b in module.buffers():
            self.comm.broadcast_flat(b, root=0)
        # bucket plan over arena slices
        self._params = list(module.parameters())
        names = self.arena.names()
        sizes = [p.numel() * 4 for p in self._params]
        self.buckets = []
        for ids in plan_buckets(sizes, first_bucket_bytes, bucket_cap_bytes):
            lo = min(self.arena.index[names[i]][0] for i in ids)
            hi = max(self.arena.index[names[i]][0] + self.arena.index[names[i]][1] for i in ids)
            self.buckets.append(dict(ids=set(ids), lo=lo, hi=hi, pending=len(ids), work=None))
        self._bucket_of = {}
        for bi, b in enumerate(self.buckets):
            for i in b["ids"]:
                self._bucket_of[i] = bi
        self._callback_queued = False
        self._comm_stream = None
        self._slots = [self.arena.index[n] for n in names]
        for i, p in enumerate(self._params):
            if p.requires_grad:
                p.register_post_accumulate_grad_hook(self._make_hook(i))

    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)

    class _NoSync:
        def __init__(self, ddp):
            self.ddp = ddp

        def __enter__(self):
            self.prev = self.ddp.require_backward_grad_sync
            self.ddp.require_backward_grad_sync = False

        def __exit__(self, *a):
            self.ddp.require_backward_grad_sync = self.prev

    def no_sync(self):
        return DistributedDataParallel._NoSync(self)

    # -- reducer --
    def _make_hook(self, i: int):
        def hook(p: torch.Tensor):
            if not self.require_backward_grad_sync or self.comm.world == 1:
                self._adopt_grad(i, p)
                return
            self._adopt_grad(i, p)
            if not self._callback_queued:
                self._callback_queued = True
                Variable._execution_engine.queue_callback(self._finalize)
            b = self.buckets[self._bucket_of[i]]
            b["pending"] -= 1
            if b["pending"] == 0:
                self._launch_bucket(b)
        return hook

    def _adopt_grad(self, i: int, p: torch.Tensor) -> None:
        """Make sure p.grad is the arena view (zero_grad(set_to_none=True) detaches it)."""
        off, k, shp = self._slots[i]
        view = self.arena.grad[off:off + k].view(shp)
        if p.grad is not None and p.grad.data_ptr() != view.data_ptr():
            view.copy_(p.grad)
            p.grad = view

    def _launch_bucket(self, b: dict) -> None:
        import torch.distributed as dist
        t = self.arena.grad[b["lo"]:b["hi"]]
        if isinstance(self.comm, TorchDistComm):
            b["work"] = dist.all_reduce(t, async_op=True)
        elif t.is_cuda:
            # in-kernel NVLink all-reduce on a side stream: it overlaps the rest of backward, which keeps
            # running on the compute stream (one-shot kernels only wait on peers' pushes, never on a local
            # grid barrier, so sharing the SMs with backward kernels cannot deadlock)
            if self._comm_stream is None:
                self._comm_stream = torch.cuda.Stream(t.device)
            ready = torch.cuda.Event()
            ready.record()
            with torch.cuda.stream(self._comm_stream):
                self._comm_stream.wait_event(ready)
                self.comm.allreduce_flat(t, average=False)
                done = torch.cuda.Event()
                done.record()
            b["work"] = done
        else:
            self.comm.allreduce_flat(t, average=False)
            b["work"] = None

    def _finalize(self) -> None:
        for b in self.buckets:
            if b["pending"] != 0:  # parameters unused in this backward: reduce what we have (zeros)
                self._launch_bucket(b)
        for b in self.buckets:
            if b["work"] is not None:
                if isinstance(b["work"], torch.cuda.Event):
                    torch.cuda.current_stream().wait_event(b["work"])
                else:
                    b["work"].wait()
                b["work"] = None
            self.arena.grad[b["lo"]:b["hi"]].div_(self.comm.world)
            b["pending"] = len(b["ids"])
        self._callback_queued = False
