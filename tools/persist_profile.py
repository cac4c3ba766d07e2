#!/usr/bin/env python
"""Per-phase clock64 stamps of one mid-sequence step of the persistent LSTM kernels (CTA 0).

Usage: python tools/persist_profile.py [--batch 64] [--hidden 1024] [--seq 512]
Prints, for forward and backward, cycle offsets relative to the epilogue's step start:
producer (grid wait begin / end / last TMA issued), MMA (first operand landed / last MMA committed),
epilogue (step start / accumulator ready / math+operand store done / CTA barrier / arrive issued).
"""
import argparse
import ctypes as C
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200 import FusedLSTM, _native  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--hidden", type=int, default=1024)
    ap.add_argument("--seq", type=int, default=512)
    a = ap.parse_args()
    lib = _native.lib()
    lib.b200rnn_persist_prof.argtypes = [C.POINTER(C.c_longlong)]
    lib.b200rnn_persist_prof.restype = C.c_int
    m = FusedLSTM(16, a.hidden, 1, batch_first=True).cuda()
    x = torch.randn(a.batch, a.seq, 16, device="cuda")
    for _ in range(2):
        out, _ = m(x)
        out[:, -1].sum().backward()
    torch.cuda.synchronize()
    buf = (C.c_longlong * 48)()
    _native.check(lib.b200rnn_persist_prof(buf), "persist_prof")
    v = list(buf)
    res = {}
    for d, name in ((0, "forward"), (1, "backward")):
        s = v[24 * d:24 * d + 24]
        base = s[16]
        res[name] = {
            "producer_wait_begin": s[0] - base, "producer_wait_end": s[1] - base, "producer_last_tma_issued": s[2] - base,
            "mma_first_operand": s[8] - base, "mma_last_commit": s[9] - base,
            "epi_acc_ready": s[17] - base, "epi_math_done": s[18] - base, "epi_cta_barrier": s[19] - base,
            "epi_arrive_issued": s[20] - base,
        }
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
