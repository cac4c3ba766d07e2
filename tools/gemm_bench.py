#!/usr/bin/env python
"""Throughput of the tcgen05 GEMM (ops.gemm.gemm_bf16) on the shapes the large-hidden RNN path uses, next to
torch.matmul in bf16 (cuBLAS) for context.  CUDA events, median of 5 after 2 warm-ups."""
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200.ops.gemm import gemm_bf16, gemm_status  # noqa: E402


def timed(fn, iters=5, warm=2):
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return sorted(ts)[len(ts) // 2]


def main():
    shapes = [("dW = dG^T x H_prev", 4096, 1024, 32768), ("gx = X x W_ih^T", 32768, 4096, 1024),
              ("dX = dG x W_ih", 32768, 1024, 4096), ("square", 8192, 8192, 8192)]
    out = []
    for name, M, N, K in shapes:
        a = torch.randn(M, K, device="cuda").bfloat16()
        b = torch.randn(N, K, device="cuda").bfloat16()
        c = torch.empty(M, N, device="cuda")
        t_own = timed(lambda: gemm_bf16(a, b, None, out=c))
        t_own_bf16 = timed(lambda: gemm_bf16(a, b, None, out_dtype=torch.bfloat16))
        t_lib = timed(lambda: torch.matmul(a, b.t()))
        fl = 2.0 * M * N * K
        out.append({"shape": name, "M": M, "N": N, "K": K, "own_fp32out_ms": t_own, "own_fp32out_tflops": fl / t_own / 1e9,
                    "own_bf16out_tflops": fl / t_own_bf16 / 1e9, "cublas_bf16_tflops": fl / t_lib / 1e9})
    print(json.dumps({"gemm_status": gemm_status(), "results": out}, indent=1))


if __name__ == "__main__":
    main()
