"""Experiment tooling (fabfile / notebook analogues): command construction, resume-by-command, analysis parsing."""
import json
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "bench"))


def test_sweep_grid_and_commands():
    import sweep
    runs = sweep.train_runs([1, 2, 4, 8], repeats=1)
    # {480,960,1440} x ranks {1,2,4,8} x {distributed, horovod} + local only at 1 rank (fabfile.py:48-66)
    assert len(runs) == 3 * (4 * 2 + 1)
    assert all(r["parameters"]["--seed"] == 123456789 and r["parameters"]["--epochs"] == 1 for r in runs)
    local = next(r for r in runs if r["trainer"] == "local")
    cmd = sweep.get_command(local, "/tmp/ds")
    assert cmd[1:3] == ["-m", "pytorch_distributed_rnn_b200"] and cmd[-1] == "local" and "--no-validation" in cmd
    dist8 = next(r for r in runs if r["trainer"] == "distributed" and r["hosts"] == 8)
    cmd = sweep.get_command(dist8, "/tmp/ds")
    assert "pytorch_distributed_rnn_b200.parallel.launch" in cmd and cmd[cmd.index("-n") + 1] == "8"


def test_sweep_resume_skips_recorded_commands(tmp_path, monkeypatch):
    import sweep
    calls = []

    class CP:
        stdout, stderr, returncode = "", "INFO:root:0: Memory Usage: 100.5, Training Duration: 1.25\nINFO:root:Training set of size 6912", 0
    monkeypatch.setattr(sweep.subprocess, "run", lambda *a, **k: calls.append(a[0]) or CP())
    runs = sweep.train_runs([1], repeats=1)[:2]
    res = tmp_path / "results.json"
    sweep.run_training(runs, "/tmp/ds", res)          # file is created if missing (Q15)
    assert len(calls) == 2 and len(json.load(open(res))["results"]) == 2
    sweep.run_training(runs, "/tmp/ds", res)          # second pass: everything already recorded
    assert len(calls) == 2
    entry = json.load(open(res))["results"][0]
    assert set(entry) >= {"command", "stdout", "stderr", "config"} and set(entry["config"]) == {"trainer", "hosts", "slots", "parameters"}


def test_analysis_parses_reference_schema_and_takes_max_over_ranks(tmp_path):
    import analysis
    res = {"results": [{"command": "c1", "stdout": "", "config": {"trainer": "distributed", "hosts": 2, "slots": 1, "parameters": {"--batch-size": 960}},
                        "stderr": "INFO:root:Training set of size 6912\nINFO:root:0: Memory Usage: 270.1, Training Duration: 100.0\nINFO:root:1: Memory Usage: 271.9, Training Duration: 106.5\n"},
                       {"command": "c1", "stdout": "", "stderr": "dup", "config": {}}]}
    f = tmp_path / "r.json"
    json.dump(res, open(f, "w"))
    rows = analysis.load([f])
    assert len(rows) == 1 and rows[0]["seconds"] == 106.5 and rows[0]["memory"] == 271.9 and rows[0]["ranks"] == 2
    assert "| distributed | 2 | 960 | 106.5000 | 65 |" in analysis.table(rows)
    ref = Path("/root/reference/evaluation/results_202007141730.json")
    if ref.exists():   # the reference's own result files parse too
        assert len(analysis.load([ref])) > 20
