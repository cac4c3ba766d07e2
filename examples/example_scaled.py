#!/usr/bin/env python
"""The motion classifier scaled up (2-layer LSTM, hidden 1024, bf16 operands) trained data-parallel with per-layer
gradient buckets — BASELINE.json's config 4 as a runnable example.  The reference has no counterpart script; its
`DistributedDataParallel(model)` (src/motion/trainer/ddp.py:19) is what overlaps bucket all-reduces with backward there.

    python examples/example_scaled.py                                  # one GPU
    python -m pytorch_distributed_rnn_b200.parallel.launch -n 8 examples/example_scaled.py

Every rank prints its per-step loss and, at the end, the overlap timeline of the last step: when each gradient bucket's
device-side trigger fired (the last weight-gradient GEMM tile of the layer retired), when its all-reduce + Adam was done,
and when the backward of each layer began and ended.
"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200 import MotionModel, ParamArena  # noqa: E402
from pytorch_distributed_rnn_b200.ops.layered_engine import LayeredTrainEngine  # noqa: E402
from pytorch_distributed_rnn_b200.parallel import process_group as pg  # noqa: E402


def main(hidden=1024, layers=2, seq_len=512, per_gpu_batch=64, steps=5):
    info = pg.init_distributed()
    torch.manual_seed(0)
    model = MotionModel(9, hidden, layers, 6).to(info.device)
    comm, arena = None, None
    if info.world_size > 1:
        from pytorch_distributed_rnn_b200.parallel.comm import make_comm
        comm = make_comm(info)
        arena = comm.make_arena(model) if hasattr(comm, "make_arena") else ParamArena(model)
        comm.broadcast_flat(arena.flat[:arena.numel], root=0)           # every rank starts from rank 0's weights
    engine = LayeredTrainEngine(model, lr=0.0025, arena=arena, comm=comm, timeline=True)
    g = torch.Generator().manual_seed(100 + info.rank)                  # every rank its own shard of synthetic data
    for step in range(steps):
        x = torch.randn(per_gpu_batch, seq_len, 9, generator=g).to(info.device)
        y = torch.randint(0, 6, (per_gpu_batch,), generator=g).to(info.device)
        m = engine.train_step(x, y).cpu()
        print(f"rank {info.rank} step {step}: loss {float(m[0]) / per_gpu_batch:.4f}  correct {int(m[1])}/{per_gpu_batch}", flush=True)
    if info.rank == 0:
        for name, b in engine.last_timeline()["buckets"].items():
            print(f"bucket {name:7s} {b['bytes'] / 2**20:7.2f} MiB  trigger {b['trigger_us']:9.1f} us  collective done "
                  f"{b['collective_done_us']:9.1f} us  backward {b['backward_begin_us']} .. {b['backward_end_us']}")
    if comm is not None and hasattr(comm, "check"):
        comm.check()
        comm.close()
    pg.shutdown()


if __name__ == "__main__":
    main()
