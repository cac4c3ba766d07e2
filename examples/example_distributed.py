#!/usr/bin/env python
"""E2 — blocking point-to-point send/recv (parity: src/example/example_distributed.py).

Rank 0 sends a 1-element tensor to every other rank; each prints ``Rank r has data 1.0``.
Launch:  python -m pytorch_distributed_rnn_b200.parallel.launch -n 2 examples/example_distributed.py
On GPUs with the symmetric-memory communicator the transfer is the K9 put+signal primitive
(a peer store followed by a release flag); otherwise torch.distributed send/recv.
"""
import torch
import torch.distributed as dist

import _path  # noqa: F401
from pytorch_distributed_rnn_b200.parallel import process_group as pg
from pytorch_distributed_rnn_b200.parallel.comm import make_comm


def run(rank, size, comm, device):
    tensor = torch.zeros(1, device=device)
    if hasattr(comm, "put"):  # K9 path
        if rank == 0:
            tensor += 1
            for i in range(1, size):
                comm.put(tensor, dst=i)
        else:
            comm.get(tensor, src=0)
    else:
        if rank == 0:
            tensor += 1
            for i in range(1, size):
                dist.send(tensor=tensor, dst=i)
        else:
            dist.recv(tensor=tensor, src=0)
    print("Rank ", rank, " has data ", float(tensor[0]))
    return float(tensor[0])


def init_process(fn):
    info = pg.init_distributed()
    comm = make_comm(info)
    out = fn(info.rank, info.world_size, comm, info.device)
    comm.barrier()
    return out


if __name__ == "__main__":
    init_process(run)
