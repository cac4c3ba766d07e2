import ctypes as C, sys, torch
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200 import _native, MotionModel
from pytorch_distributed_rnn_b200.ops.engine import FusedTrainEngine
lib=_native.lib(); lib.b200rnn_tc_prof.argtypes=[C.c_void_p]; lib.b200rnn_tc_prof.restype=C.c_int
for B in (100, 180, 1440, 4000):
    m=MotionModel(9,32,2,6).cuda(); e=FusedTrainEngine(m); x=torch.randn(B,128,9,device='cuda'); buf=e._get(B,128); d=buf['desc']
    for _ in range(3): lib.b200rnn_lstm_fwd_tc(C.byref(d), e.arena.flat.data_ptr(), x.data_ptr(), buf['stash'].data_ptr(), 1, _native.stream_ptr())
    torch.cuda.synchronize()
    out=(C.c_longlong*16)(); lib.b200rnn_tc_prof(out)
    for who in (0,1):
        st=[out[who*8+i] for i in range(8)]
        print(B, "tid0" if who==0 else "tid127", [st[i+1]-st[i] for i in range(7)], "total", st[7]-st[0])
