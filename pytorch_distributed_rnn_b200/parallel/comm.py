"""Gradient-path communicators.

Two implementations of the same small surface (``allreduce_flat``, ``broadcast_flat``,
``allreduce_update``, ``barrier``):

* :class:`TorchDistComm` — ``torch.distributed`` collectives (gloo on CPU, NCCL on GPU).
  This is the *baseline*: what the reference gets from c10d/MPI
  (src/motion/trainer/ddp.py:18-19) and what stock DDP would do on B200.
* :class:`~.symm.SymmComm` — the product path on NVLink: peer-mapped symmetric buffers and
  a hand-written kernel that does reduction of the per-CTA gradient rows, one-shot /
  two-shot all-reduce by P2P stores, gradient averaging and the optimizer step in one
  launch (csrc/comm).

``make_comm`` picks SymmComm on multi-GPU CUDA worlds unless ``B200RNN_COMM=nccl``.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch
import torch.distributed as dist

from .. import _native
from .process_group import RankInfo, rank_info


class TorchDistComm:
    kind = "torch"

    def __init__(self, info: Optional[RankInfo] = None):
        self.info = info or rank_info()
        self.rank, self.world, self.device = self.info.rank, self.info.world_size, self.info.device

    def allreduce_flat(self, t: torch.Tensor, average: bool = True) -> torch.Tensor:
        if self.world > 1:
            dist.all_reduce(t)
            if average:
                t.div_(self.world)
        return t

    def broadcast_flat(self, t: torch.Tensor, root: int = 0) -> torch.Tensor:
        if self.world > 1:
            dist.broadcast(t, src=root)
        return t

    def barrier(self) -> None:
        if self.world > 1:
            dist.barrier()

    # fused-engine hook: rows -> local reduce -> NCCL all-reduce -> optimizer (3 launches, baseline)
    def allreduce_update(self, rows: torch.Tensor, n_rows: int, arena, opt, optdesc, metrics_slot: torch.Tensor) -> None:
        lib = _native.lib()
        sp = _native.stream_ptr(self.device)
        row = arena.numel + 2
        _native.check(lib.b200rnn_reduce_rows(rows.data_ptr(), n_rows, row, arena.grad.data_ptr(), 0, sp), "reduce_rows")
        metrics_slot.copy_(arena.grad[arena.numel:arena.numel + 2])  # this rank's loss sum / correct count
        dist.all_reduce(arena.grad[:arena.numel])
        _native.check(lib.b200rnn_fused_update(arena.grad.data_ptr(), 1, arena.numel, arena.flat.data_ptr(),
                                               opt.exp_avg.data_ptr(), opt.exp_avg_sq.data_ptr(), C.byref(optdesc),
                                               None, None, sp), "fused_update")
        _native.count_launch(2)


def make_comm(info: Optional[RankInfo] = None, prefer: Optional[str] = None):
    info = info or rank_info()
    prefer = prefer or os.environ.get("B200RNN_COMM", "auto")
    if info.world_size > 1 and info.device.type == "cuda" and prefer in ("auto", "symm"):
        comm, err = None, None
        try:
            from .symm import SymmComm
            comm = SymmComm(info)
        except Exception as e:  # noqa: BLE001 - report and fall back to the NCCL baseline only when allowed
            err = e
        # every rank must end up with the same kind of communicator: a rank that fell back to NCCL while its peers
        # spin in the NVLink kernels would hang the job, so the fallback is taken by all ranks or by none
        ok = torch.tensor([1 if comm is not None else 0], device=info.device)
        if dist.is_initialized():
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)
        if int(ok.item()) == 1:
            return comm
        if comm is not None:
            comm.close()
        if prefer == "symm":
            raise RuntimeError(f"symmetric-memory communicator unavailable on this or a peer rank: {err}")
        import logging
        logging.warning("symmetric-memory communicator unavailable (%s); all ranks use the NCCL baseline",
                        err if err is not None else "a peer rank failed to set it up")
    return TorchDistComm(info)
