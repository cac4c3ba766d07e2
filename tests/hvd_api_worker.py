"""World-size-2 worker for tests/test_distributed_cpu.py::test_horovod_style_api_surface (gloo on CPU)."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200.parallel import horovod as hvd  # noqa: E402

hvd.init()
r, w = hvd.rank(), hvd.size()
assert hvd.is_initialized() and hvd.local_size() == w and hvd.local_rank() == r
x = torch.arange(6, dtype=torch.float32) + r
assert torch.equal(hvd.allreduce(x), torch.arange(6, dtype=torch.float32) + (w - 1) / 2)          # average
assert torch.equal(hvd.allreduce(x, op=hvd.Sum), w * torch.arange(6, dtype=torch.float32) + w * (w - 1) / 2)
assert torch.equal(x, torch.arange(6, dtype=torch.float32) + r)                                     # not in place
y = x.clone()
assert hvd.allreduce_(y, average=False) is y and torch.equal(y, hvd.allreduce(x, op=hvd.Sum))
try:
    hvd.allreduce(x, average=True, op=hvd.Sum)
    raise SystemExit("average= and op= together must be refused")
except ValueError:
    pass
b = hvd.broadcast(torch.full((2, 3), float(r)), root_rank=1)
assert torch.equal(b, torch.ones(2, 3))
z = torch.full((4,), float(r))
hvd.broadcast_(z, 0)
assert torch.equal(z, torch.zeros(4))
g = hvd.allgather(torch.full((r + 1, 2), float(r)))            # uneven first dimension
assert g.shape == (sum(range(1, w + 1)), 2) and torch.equal(g[:1], torch.zeros(1, 2)) and torch.equal(g[1:3], torch.ones(2, 2))
hvd.join()
hvd.shutdown()
assert not hvd.is_initialized()
print(f"rank {r}: hvd api ok")
