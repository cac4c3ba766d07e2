#!/usr/bin/env python
"""Three bulk GEMMs (8192^3, bf16 operands, fp32 output) on the CTA-pair persistent kernel, for ncu:
ncu --set full --clock-control none --import-source on -k regex:gemm_bf16_pair -s 1 -c 1 python tools/gemm_ncu_target.py"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200.ops.gemm import gemm_bf16, gemm_status  # noqa: E402

M = N = K = 8192
a = torch.randn(M, K, device="cuda").bfloat16()
b = torch.randn(N, K, device="cuda").bfloat16()
c = torch.empty(M, N, device="cuda")
for _ in range(3):
    gemm_bf16(a, b, None, out=c)
torch.cuda.synchronize()
assert gemm_status() == 0
print("ok", float(c[0, 0]))
