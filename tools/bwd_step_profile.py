"""clock64 breakdown of one backward step (middle of the sequence): the CTA of block 0 (a layer that also produces
dx: the top layer in wave mode) and the last CTA (bottom layer).  Role B is the serial chain (recurrent product ->
gate gradients), role A accumulates dW one step behind."""
import ctypes as C, sys, torch
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200 import _native, MotionModel
from pytorch_distributed_rnn_b200.ops.engine import FusedTrainEngine
lib = _native.lib(); lib.b200rnn_bwd_prof.argtypes = [C.c_void_p]; lib.b200rnn_bwd_prof.restype = C.c_int
for B in (180, 360, 720, 1440):
    m = MotionModel(9, 32, 2, 6).cuda(); e = FusedTrainEngine(m)
    x = torch.randn(B, 128, 9, device='cuda'); y = torch.randint(0, 6, (B,), device='cuda')
    for _ in range(3): e.train_step(x, y)
    torch.cuda.synchronize()
    out = (C.c_longlong * 32)(); lib.b200rnn_bwd_prof(out)
    r = [[out[8 * k + i] for i in range(8)] for k in range(4)]
    buf = e._get(B, 128)
    print(B, "tile", buf["rows_per_cta"], "wave" if buf["bwd_wave"] else "seq")
    for name, a, b in (("dx-layer", r[0], r[1]), ("bottom  ", r[2], r[3])):
        print("   %s roleA: dW+rearm %d | stage %d | barrier %d | period %d" % (name, a[1]-a[0], a[2]-a[1], a[3]-a[2], a[4]-a[0]),
              "|| roleB: product %d | gate grads %d | barrier %d | period %d" % (b[1]-b[0], b[2]-b[1], b[3]-b[2], b[4]-b[0]))
