#!/usr/bin/env python
"""Two fused training steps of the motion model at a given per-rank batch (default 180 = 8-GPU share), for ncu:
ncu --set full -k regex:lstm_(fwd|bwd)_kernel -c 2 python tools/small_step_once.py [--batch 180] [--cell lstm]"""
import argparse
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200 import MotionModel  # noqa: E402
from pytorch_distributed_rnn_b200.ops.engine import FusedTrainEngine  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=180)
ap.add_argument("--cell", default="lstm")
a = ap.parse_args()
torch.manual_seed(0)
model = MotionModel(9, 32, 2, 6, cell=a.cell).cuda()
eng = FusedTrainEngine(model, lr=0.0025)
x = torch.randn(a.batch, 128, 9, device="cuda")
y = torch.randint(0, 6, (a.batch, 1), device="cuda")
for _ in range(2):
    m = eng.train_step(x, y)
torch.cuda.synchronize()
print("ok", m.cpu().tolist())
