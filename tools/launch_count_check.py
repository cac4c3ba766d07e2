#!/usr/bin/env python
"""Print how many kernels one fused training step / one evaluation launches at a few per-rank batch sizes
(3 per step; +1 per layer-wavefront kernel for its sentinel pre-fill).  bench.py reports the same counter."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200 import MotionModel, _native  # noqa: E402
from pytorch_distributed_rnn_b200.ops.engine import FusedTrainEngine  # noqa: E402

m = MotionModel(9, 32, 2, 6).cuda()
e = FusedTrainEngine(m)
for B in (180, 360, 720, 1440):
    x, y = torch.randn(B, 128, 9, device="cuda"), torch.randint(0, 6, (B, 1), device="cuda")
    b = _native.launches(); e.train_step(x, y); s = _native.launches() - b
    b = _native.launches(); e.evaluate(x, y); v = _native.launches() - b
    print(f"B={B}: train step {s} launches, evaluate {v}")
torch.cuda.synchronize()
print(_native.device_status())
