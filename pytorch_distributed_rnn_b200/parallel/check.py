"""Connectivity / topology smoke test — the analogue of the reference's cluster preparation.

The reference checks its Pi cluster with ``ssh-keyscan`` + ``mpirun --host ... hostname``
(fabfile.py:69-77, README step 8) before running anything.  On an NVSwitch node the equivalent
questions are: do all ranks rendezvous, can every GPU address every peer (NVLink P2P, NVLS
multicast support), and do a barrier / broadcast / all-reduce over the in-kernel path and over NCCL
agree?  Run under the launcher:

    python -m pytorch_distributed_rnn_b200.parallel.launch -n 8 -m pytorch_distributed_rnn_b200.parallel.check
"""
from __future__ import annotations

import json
import socket

import torch
import torch.distributed as dist

from . import process_group as pg
from .comm import TorchDistComm, make_comm


def main() -> int:
    info = pg.init_distributed()
    rank, world, dev = info.rank, info.world_size, info.device
    report = {"rank": rank, "world": world, "host": socket.gethostname(), "device": str(dev), "backend": info.backend}
    comm = make_comm(info)
    report["comm"] = comm.kind
    if dev.type == "cuda":
        from .bootstrap import probe_topology
        t = probe_topology()
        report["gpu"] = t["names"][dev.index]
        report["all_to_all_p2p"] = t["all_to_all"]
        report["nvls_multicast"] = t["multicast_supported"][dev.index]
        if rank == 0:
            print("P2P access matrix:", json.dumps(t["can_access"]))
    v = torch.full((1024,), float(rank + 1), device=dev)
    comm.allreduce_flat(v, average=False)
    ref = torch.full((1024,), float(rank + 1), device=dev)
    TorchDistComm(info).allreduce_flat(ref, average=False)
    b = torch.full((16,), float(rank), device=dev)
    comm.broadcast_flat(b, root=0)
    comm.barrier()
    if hasattr(comm, "check"):
        comm.check()
    ok = bool(torch.equal(v, ref)) and float(v[0]) == world * (world + 1) / 2 and float(b[0]) == 0.0
    report["ok"] = ok
    print(json.dumps(report), flush=True)
    if world > 1:
        dist.barrier()
    if hasattr(comm, "close"):
        comm.close()
    pg.shutdown()
    return 0 if ok else 1


if __name__ == "__main__":
    raise SystemExit(main())
