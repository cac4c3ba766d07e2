"""Process-group bootstrap: one process per GPU, rendezvous from the launcher's environment.

Replaces the reference's ``init_process_group('mpi')`` under ``mpirun``
(src/motion/trainer/ddp.py:18, src/example/example_ddp.py:98) and ``hvd.init()`` under
``horovodrun`` (trainer/horovod.py:17).  Rank / world size are read from, in order:
``torchrun`` variables (``RANK``/``WORLD_SIZE``/``LOCAL_RANK``/``MASTER_ADDR``/``MASTER_PORT``),
this package's own launcher (:mod:`.launch`, same variables), or OpenMPI's
``OMPI_COMM_WORLD_*`` so an existing ``mpirun`` command line keeps working.  With none of
them set the process is a world of one.  Backend: NCCL when CUDA is available (it carries
only the plumbing collectives — the gradient path is the in-kernel NVLink all-reduce),
gloo on CPU.
"""
from __future__ import annotations

import datetime
import os
import tempfile
from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist


@dataclass
class RankInfo:
    rank: int
    world_size: int
    local_rank: int
    device: torch.device
    backend: str


def env_rank_info() -> tuple[int, int, int]:
    e = os.environ
    if "RANK" in e and "WORLD_SIZE" in e:
        return int(e["RANK"]), int(e["WORLD_SIZE"]), int(e.get("LOCAL_RANK", e["RANK"]))
    if "OMPI_COMM_WORLD_RANK" in e:
        return (int(e["OMPI_COMM_WORLD_RANK"]), int(e["OMPI_COMM_WORLD_SIZE"]),
                int(e.get("OMPI_COMM_WORLD_LOCAL_RANK", e["OMPI_COMM_WORLD_RANK"])))
    return 0, 1, 0


_info: Optional[RankInfo] = None


def init_distributed(backend: Optional[str] = None, use_cuda: Optional[bool] = None,
                     timeout_s: int = 300) -> RankInfo:
    """Idempotent. Initialises torch.distributed (if not already) and selects this rank's GPU."""
    global _info
    if _info is not None and dist.is_initialized():
        return _info
    rank, world, local = env_rank_info()
    cuda = torch.cuda.is_available() if use_cuda is None else (use_cuda and torch.cuda.is_available())
    if cuda:
        ndev = torch.cuda.device_count()
        device = torch.device("cuda", local % max(ndev, 1))
        torch.cuda.set_device(device)
    else:
        device = torch.device("cpu")
    backend = backend or ("nccl" if cuda else "gloo")
    if not dist.is_initialized():
        kw = dict(backend=backend, rank=rank, world_size=world, timeout=datetime.timedelta(seconds=timeout_s))
        if "MASTER_ADDR" in os.environ and "MASTER_PORT" in os.environ:
            dist.init_process_group(init_method="env://", **kw)
        elif world == 1:
            store = os.path.join(tempfile.mkdtemp(prefix="b200rnn_pg_"), "store")
            dist.init_process_group(init_method=f"file://{store}", **kw)
        else:
            raise RuntimeError("world_size > 1 needs MASTER_ADDR / MASTER_PORT (use torchrun or "
                               "python -m pytorch_distributed_rnn_b200.parallel.launch)")
    _info = RankInfo(dist.get_rank(), dist.get_world_size(), local, device, backend)
    return _info


def rank_info() -> RankInfo:
    return _info if _info is not None else init_distributed()


def shutdown() -> None:
    global _info
    if dist.is_initialized():
        dist.destroy_process_group()
    _info = None
