"""End-to-end trainer on the GPU: same log lines / numbers as the CPU torch path, checkpoints load into stock torch."""
import logging
import re

import pytest
import torch
from torch import nn

from pytorch_distributed_rnn_b200 import MotionDataset, MotionModel
from pytorch_distributed_rnn_b200.data import synthetic
from pytorch_distributed_rnn_b200.trainer import Trainer

pytestmark = pytest.mark.gpu
STEP_RE = re.compile(r"Train Batch: (\d+)/(\d+) .*Loss: ([\d.]+)\tAcc: (\d+)/(\d+)")


def _run(device, caplog, tmp_path, deferred=False):
    (xtr, ytr), (xv, yv), (xte, yte) = synthetic.make_datasets(seed=0, train=960, valid=50, test=60)
    torch.manual_seed(123456789)
    model = MotionModel(9, 32, 2, 6)
    t = Trainer(model, MotionDataset(xtr, ytr), 480, 0.0025, validation_set=MotionDataset(xv, yv),
                test_set=MotionDataset(xte, yte), checkpoint_dir=tmp_path / f"ck_{device}", device=device,
                deferred_logging=deferred)
    caplog.clear()
    with caplog.at_level(logging.INFO):
        _, th, vh = t.train(2)
    steps = [(int(m[1]), float(m[3]), int(m[4])) for m in STEP_RE.finditer(caplog.text)]
    return t, th, vh, steps, caplog.text


def test_gpu_trainer_matches_cpu_trainer(caplog, tmp_path):
    tc, thc, vhc, sc, _ = _run("cpu", caplog, tmp_path)
    tg, thg, vhg, sg, text = _run("cuda", caplog, tmp_path)
    assert tg.engine is not None, "the fused engine must be the path that runs on CUDA"
    assert len(sc) == len(sg) == 4
    for (b1, l1, c1), (b2, l2, c2) in zip(sc, sg):
        assert b1 == b2 and abs(l1 - l2) < 2e-5 and abs(c1 - c2) <= 1
    assert all(abs(a - b) < 1e-6 for a, b in zip(thc, thg)) and all(abs(a - b) < 1e-4 for a, b in zip(vhc, vhg))
    assert re.search(r"0: Memory Usage: \d+\.\d+, Training Duration: \d+\.\d+", text)
    assert "Test Evaluation:" in text
    ck = torch.load(tmp_path / "ck_cuda" / "best-model.pt", weights_only=False)
    ref = nn.LSTM(9, 32, 2, batch_first=True)
    ref.load_state_dict({k[5:]: v for k, v in ck["model_state"].items() if k.startswith("lstm.")})
    assert ck["optimizer_state"]["state"][0]["exp_avg"].shape == (128, 9)


def test_deferred_logging_same_lines(caplog, tmp_path):
    _, _, _, s1, _ = _run("cuda", caplog, tmp_path)
    _, _, _, s2, _ = _run("cuda", caplog, tmp_path, deferred=True)
    assert s1 == s2


def test_pipelined_host_steps_match_serial_steps():
    from pytorch_distributed_rnn_b200.ops.engine import FusedTrainEngine
    xs = [torch.randn(96, 32, 9).pin_memory() for _ in range(5)]
    ys = [torch.randint(0, 6, (96,)).pin_memory() for _ in range(5)]
    torch.manual_seed(3)
    a = FusedTrainEngine(MotionModel(9, 32, 2, 6).cuda())
    torch.manual_seed(3)
    b = FusedTrainEngine(MotionModel(9, 32, 2, 6).cuda())
    serial = [a.train_step_host(x, y) for x, y in zip(xs, ys)]
    piped = list(b.train_steps_host(zip(xs, ys)))
    assert serial == piped
    assert torch.equal(a.arena.flat, b.arena.flat)


def test_pinned_dataset_steps_match_device_resident_steps():
    """train_steps_dataset (index slice -> device, rows pulled from the pinned host dataset by the gather kernel)
    must train exactly like stepping through the same index slices of a device-resident copy."""
    from pytorch_distributed_rnn_b200.ops.engine import FusedTrainEngine
    torch.manual_seed(5)
    N, B = 500, 96
    feats, labels = torch.randn(N, 32, 9), torch.randint(0, 6, (N, 1))
    idx = [torch.randperm(N)[:B] for _ in range(5)] + [torch.randperm(N)[:40]]      # last batch partial
    torch.manual_seed(3)
    a = FusedTrainEngine(MotionModel(9, 32, 2, 6).cuda())
    torch.manual_seed(3)
    b = FusedTrainEngine(MotionModel(9, 32, 2, 6).cuda())
    fd, ld = feats.cuda(), labels.cuda()
    want = []
    for i in idx:
        m = a.train_step_gather(fd, ld, i.cuda()).cpu()
        want.append((float(m[0]) / i.numel(), int(round(float(m[1])))))
    got = list(b.train_steps_dataset(feats.pin_memory(), labels.pin_memory(), idx))
    assert got == want
    assert torch.equal(a.arena.flat, b.arena.flat)
