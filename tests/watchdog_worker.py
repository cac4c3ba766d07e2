"""Worker of test_watchdog_trips_instead_of_hanging: the last rank skips the third all-reduce."""
import os
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200.parallel import process_group as pg  # noqa: E402
from pytorch_distributed_rnn_b200.parallel.symm import SymmComm  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    info = pg.init_distributed()
    rank, world, dev = info.rank, info.world_size, info.device
    comm = SymmComm(info, inbox_floats=1 << 20, heap_bytes=32 << 20)
    v = torch.ones(14150, device=dev)
    for _ in range(2):
        comm.allreduce_flat(v, average=True)
    torch.cuda.synchronize()
    comm.check()
    dist.barrier()
    code = 0
    if rank != world - 1:
        comm.allreduce_flat(v, average=True)      # the last rank never joins this one
        torch.cuda.synchronize()
        try:
            comm.check()
            print("WATCHDOG_MISSED", flush=True)
            code = 4
        except RuntimeError as e:
            print(f"WATCHDOG_TRIPPED {e}", flush=True)
            code = 3
    sys.stdout.flush()
    os._exit(code)      # no collective teardown: the peers are in different states on purpose


if __name__ == "__main__":
    main()
