"""Multi-process plumbing on CPU (gloo, world_size 2-3): the reference's only correctness checks.

* toy DDP / Horovod-style entry points end with identical parameters on every rank
  (README.md:9, src/example/example_ddp.py:83-87);
* world-size invariance of the per-step mean loss and summed correct count;
* DDP checkpoints carry the ``module.`` prefix, Horovod/local ones do not (Q9);
* the asynchronous parameter server trains and shuts down cleanly.
"""
import json
import re
import sys
from pathlib import Path

import pytest
import torch

from pytorch_distributed_rnn_b200.parallel.launch import launch

ROOT = Path(__file__).resolve().parent.parent
PY = sys.executable
STEP_RE = re.compile(r"Rank: (\d+)   Train Batch: (\d+)/(\d+) .*Loss: ([\d.]+)\tAcc: (\d+)/(\d+)")


def _ok(results):
    for r, cp in enumerate(results):
        assert cp.returncode == 0, f"rank {r} failed:\n{cp.stdout}\n{cp.stderr}"


def _final_params(results):
    vals = []
    for cp in results:
        lines = [l for l in cp.stdout.splitlines() if "parameters:" in l]
        vals.append(float(lines[-1].split("parameters:")[1]))
    return vals


@pytest.mark.parametrize("script,extra", [("example_ddp.py", []), ("example_ddp.py", ["--sharded"]), ("example_horovod.py", [])])
def test_toy_examples_identical_final_parameters(script, extra):
    res = launch([PY, str(ROOT / "examples" / script), *extra], 2, timeout=240, env={"OMP_NUM_THREADS": "1"})
    _ok(res)
    vals = _final_params(res)
    assert vals[0] == vals[1], vals  # bit-identical, not merely close


def test_p2p_example():
    res = launch([PY, str(ROOT / "examples" / "example_distributed.py")], 3, timeout=240, env={"OMP_NUM_THREADS": "1"})
    _ok(res)
    for r, cp in enumerate(res):
        assert f"Rank  {r}  has data  1.0" in cp.stdout


def _run_cli(nproc, trainer, ds, tmp, extra=()):
    cmd = [PY, "-m", "pytorch_distributed_rnn_b200", "--dataset-path", str(ds), "--epochs", "1", "--batch-size", "96",
           "--seed", "123456789", "--device", "cpu", "--checkpoint-directory", str(tmp / f"ck_{trainer}_{nproc}"),
           "--history-path", str(tmp / f"h_{trainer}_{nproc}.json"), *extra, trainer]
    res = launch(cmd, nproc, timeout=600, env={"OMP_NUM_THREADS": "1", "PYTHONPATH": str(ROOT)})
    _ok(res)
    steps = {}
    for cp in res:
        for m in STEP_RE.finditer(cp.stderr):
            rank, b, nb, loss, corr, n = int(m[1]), int(m[2]), int(m[3]), float(m[4]), int(m[5]), int(m[6])
            steps.setdefault(b, []).append((rank, loss, corr, n))
    return res, steps


def test_world_size_invariance_and_checkpoint_prefix(tiny_dataset_dir, tmp_path):
    _, local = _run_cli(1, "local", tiny_dataset_dir, tmp_path, ["--no-validation"])
    for trainer in ("distributed", "horovod"):
        _, dist2 = _run_cli(2, trainer, tiny_dataset_dir, tmp_path, ["--no-validation"])
        assert sorted(dist2) == sorted(local)
        for b in local:
            (_, l_loss, l_corr, l_n), = local[b]
            assert len(dist2[b]) == 2
            mean = sum(x[1] for x in dist2[b]) / 2
            assert abs(mean - l_loss) < 5e-6, (trainer, b, mean, l_loss)
            assert sum(x[2] for x in dist2[b]) == l_corr
            assert sum(x[3] for x in dist2[b]) == l_n
    # with validation: checkpoints + key prefixes; history.json written once, by rank 0
    res, _ = _run_cli(2, "distributed", tiny_dataset_dir, tmp_path)
    ck = torch.load(tmp_path / "ck_distributed_2" / "best-model.pt", weights_only=False)
    assert all(k.startswith("module.") for k in ck["model_state"]) and ck["epoch"] == 1
    assert "0: Memory Usage:" in res[0].stderr and "1: Memory Usage:" in res[1].stderr
    assert "Evaluation Epoch: 1/1" in res[0].stderr and "Evaluation Epoch" not in res[1].stderr
    res, _ = _run_cli(2, "horovod", tiny_dataset_dir, tmp_path)
    ck = torch.load(tmp_path / "ck_horovod_2" / "best-model.pt", weights_only=False)
    assert not any(k.startswith("module.") for k in ck["model_state"])
    hist = json.load(open(tmp_path / "h_horovod_2.json"))
    assert set(hist) == {"train_history", "validation_history"} and len(hist["train_history"]) == 1


def test_parameter_server_cpu(tiny_dataset_dir, tmp_path):
    import socket
    import subprocess
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    procs = []
    for rank in range(3):
        cmd = [PY, "-m", "pytorch_distributed_rnn_b200", "--dataset-path", str(tiny_dataset_dir), "--epochs", "1",
               "--batch-size", "32", "--seed", "1", "--device", "cpu", "--num-threads", "1", "--no-validation",
               "parameter-server", "--world-size", "3", "--rank", str(rank),
               "--master-address", "127.0.0.1", "--master-port", str(port)]
        procs.append(subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True,
                                      env={**__import__("os").environ, "PYTHONPATH": str(ROOT), "OMP_NUM_THREADS": "1"}))
    outs = []
    try:
        outs = [p.communicate(timeout=180) for p in procs]
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    for p, (o, e) in zip(procs, outs):
        assert p.returncode == 0, o + e
    assert "RPC shutdown on parameter server. (4 asynchronous updates applied)" in outs[0][0]  # 2 workers x 64/32 batches
    for r in (1, 2):
        assert len(STEP_RE.findall(outs[r][1])) == 2 and f"{r}: Memory Usage:" in outs[r][1]


def test_horovod_style_api_surface():
    """The everyday horovod.torch vocabulary beyond what the reference's trainer touches (allreduce ops, in-place
    forms, broadcast, uneven allgather, join, shutdown) on gloo, world size 2."""
    res = launch([PY, str(ROOT / "tests" / "hvd_api_worker.py")], 2, timeout=240, env={"OMP_NUM_THREADS": "1"})
    _ok(res)
    assert all("hvd api ok" in cp.stdout for cp in res)
