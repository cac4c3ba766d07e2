"""clock64 breakdown of one backward step (layer 1, middle of the sequence, CTA 0)."""
import ctypes as C, sys, torch
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200 import _native, MotionModel
from pytorch_distributed_rnn_b200.ops.engine import FusedTrainEngine
lib = _native.lib(); lib.b200rnn_bwd_prof.argtypes = [C.c_void_p]; lib.b200rnn_bwd_prof.restype = C.c_int
for B in (180, 720, 1440):
    m = MotionModel(9, 32, 2, 6).cuda(); e = FusedTrainEngine(m)
    x = torch.randn(B, 128, 9, device='cuda'); y = torch.randint(0, 6, (B,), device='cuda')
    for _ in range(3): e.train_step(x, y)
    torch.cuda.synchronize()
    out = (C.c_longlong * 16)(); lib.b200rnn_bwd_prof(out)
    a = [out[i] for i in range(8)]; b = [out[8 + i] for i in range(8)]
    print(B, "tile", e._get(B, 128)["rows_per_cta"],
          "roleA: phaseA %d | wait S1 %d | B2 %d | wait S2 %d" % (a[1]-a[0], a[2]-a[1], a[3]-a[2], a[4]-a[3]),
          "|| roleB: issue loads %d | wait S1 %d | B1 %d | stage %d | wait S2 %d" % (b[1]-b[0], b[2]-b[1], b[3]-b[2], b[4]-b[3], b[5]-b[4]),
          "|| step total", a[4]-a[0])
