#!/usr/bin/env python
"""Per-kernel device timings of the fused training step (CUDA events, L2 flushed between iterations)."""
import argparse
import ctypes as C
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200 import MotionModel, _native  # noqa: E402
from pytorch_distributed_rnn_b200.ops.engine import FusedTrainEngine  # noqa: E402


def time_it(fn, flush, iters=20, warm=3):
    lib = _native.lib()
    sp = _native.stream_ptr()
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        lib.b200rnn_l2_flush(flush.data_ptr(), flush.numel(), sp)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", default="1440,720,360,180")
    ap.add_argument("--T", type=int, default=128)
    a = ap.parse_args()
    dev = torch.device("cuda", 0)
    lib = _native.lib()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    out = {}
    for B in [int(x) for x in a.batches.split(",")]:
        torch.manual_seed(0)
        model = MotionModel(9, 32, 2, 6).to(dev)
        eng = FusedTrainEngine(model)
        x = torch.randn(B, a.T, 9, device=dev)
        y = torch.randint(0, 6, (B,), device=dev)
        buf = eng._get(B, a.T)
        d, ar, opt = buf["desc"], eng.arena, eng.optimizer
        sp = _native.stream_ptr()
        opt.step_count = 1
        o = opt.desc()
        fwd = lambda: lib.b200rnn_lstm_fwd(C.byref(d), ar.flat.data_ptr(), x.data_ptr(), buf["stash"].data_ptr(), 1, sp)
        fwd_tc = lambda: lib.b200rnn_lstm_fwd_tc(C.byref(d), ar.flat.data_ptr(), x.data_ptr(), buf["stash"].data_ptr(), 1, sp)
        bwd = lambda: lib.b200rnn_lstm_bwd(C.byref(d), ar.flat.data_ptr(), x.data_ptr(), buf["stash"].data_ptr(), y.data_ptr(),
                                           1.0 / B, None, None, None, buf["scratch"].data_ptr(), sp)
        upd = lambda: lib.b200rnn_fused_update(buf["scratch"].data_ptr(), buf["n_cta"], ar.numel, ar.flat.data_ptr(),
                                               opt.exp_avg.data_ptr(), opt.exp_avg_sq.data_ptr(), C.byref(o), None, None, sp)
        step = lambda: eng.train_step(x, y)
        out[B] = {"fwd_us": time_it(fwd, flush), "fwd_tc_us": time_it(fwd_tc, flush), "bwd_us": time_it(bwd, flush), "update_us": time_it(upd, flush),
                  "step_us": time_it(step, flush), "n_cta_bwd": buf["n_cta"]}
        print(B, json.dumps(out[B]), flush=True)


if __name__ == "__main__":
    main()
