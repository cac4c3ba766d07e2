#!/usr/bin/env python
"""One warm-up and one measured forward+backward of the scaled configuration (ours only), for launch lists:
ncu --metrics gpu__time_duration.sum --csv python tools/scaled_step_once.py [--cell gru] [--batch 64]"""
import argparse
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200 import FusedGRU, FusedLSTM  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--cell", default="lstm")
ap.add_argument("--batch", type=int, default=64)
a = ap.parse_args()
m = (FusedLSTM if a.cell == "lstm" else FusedGRU)(9, 1024, 2, batch_first=True).cuda()
x = torch.randn(a.batch, 512, 9, device="cuda")
for _ in range(2):
    out, _ = m(x)
    out[:, -1].sum().backward()
torch.cuda.synchronize()
print("ok")
