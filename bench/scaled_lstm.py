#!/usr/bin/env python
"""Scaled-up configuration of BASELINE.json: 2-layer LSTM, hidden 1024, seq_len 512, bf16 operands.

Times forward+backward of ``FusedLSTM`` (large-hidden path: tcgen05 GEMMs + fused cell kernels) on
one GPU with CUDA events, and — for context only — the same step of stock ``torch.nn.LSTM`` under
bf16 autocast (cuDNN).  Usage: python bench/scaled_lstm.py [--batch 64] [--hidden 1024] [--seq 512]
"""
import argparse
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200 import FusedGRU, FusedLSTM  # noqa: E402


def timed(fn, iters=3, warm=1):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--hidden", type=int, default=1024)
    ap.add_argument("--seq", type=int, default=512)
    ap.add_argument("--input", type=int, default=9)
    ap.add_argument("--layers", type=int, default=2)
    ap.add_argument("--cell", default="lstm", choices=["lstm", "gru"])
    a = ap.parse_args()
    torch.manual_seed(0)
    x = torch.randn(a.batch, a.seq, a.input, device="cuda")
    ours = (FusedLSTM if a.cell == "lstm" else FusedGRU)(a.input, a.hidden, a.layers, batch_first=True).cuda()
    stock = (torch.nn.LSTM if a.cell == "lstm" else torch.nn.GRU)(a.input, a.hidden, a.layers, batch_first=True).cuda()
    stock.load_state_dict(ours.state_dict())

    def step_ours():
        out, _ = ours(x)
        out[:, -1].sum().backward()

    def step_stock():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out, _ = stock(x)
        out[:, -1].float().sum().backward()

    flops = 0
    for l in range(a.layers):
        i = a.input if l == 0 else a.hidden
        flops += 2 * (4 if a.cell == "lstm" else 3) * a.hidden * (i + a.hidden) * a.seq * a.batch * 3   # fwd + 2x bwd
    def fwd_ours():
        with torch.no_grad():
            ours(x)

    def fwd_stock():
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            stock(x)

    t_ours, t_stock = timed(step_ours), timed(step_stock)
    t_fwd_ours, t_fwd_stock = timed(fwd_ours), timed(fwd_stock)
    from pytorch_distributed_rnn_b200.ops.gemm import gemm_status, persist_status
    with torch.no_grad():   # sanity: same weights, so the two outputs agree to bf16 rounding
        o1 = ours(x)[0].float()
        o2 = stock(x)[0].float()
        out_rel_err = float((o1 - o2).abs().max() / (o2.abs().max() + 1e-6))
    print(json.dumps({"config": vars(a), "ours_ms": t_ours, "stock_cudnn_bf16_ms": t_stock,
                      "ours_fwd_ms": t_fwd_ours, "stock_fwd_ms": t_fwd_stock,
                      "persist_env": __import__("os").environ.get("B200RNN_PERSIST", "1"),
                      "gemm_status": gemm_status(), "persist_status": persist_status(), "out_rel_err_vs_fp32": out_rel_err,
                      "ours_tflops": flops / t_ours / 1e9, "stock_tflops": flops / t_stock / 1e9,
                      "samples_per_s_ours": a.batch / t_ours * 1e3}))


if __name__ == "__main__":
    main()
