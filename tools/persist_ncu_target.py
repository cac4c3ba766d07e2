#!/usr/bin/env python
"""Small driver for an ncu capture of the persistent LSTM kernels: one layer, H=1024, B=64, T=128, one fwd+bwd."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200 import FusedLSTM  # noqa: E402

m = FusedLSTM(16, 1024, 1, batch_first=True).cuda()
x = torch.randn(64, 128, 16, device="cuda")
out, _ = m(x)
out[:, -1].sum().backward()
torch.cuda.synchronize()
print("ok")
