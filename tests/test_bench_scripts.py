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


def test_network_log_split_and_loss_rule_tables(tmp_path):
    """O2 (fabfile.py:146-191): one sweep job's log becomes one result entry per (trainer, rule); analysis tabulates
    delay and loss rules separately, as Experiments_network.ipynb plots them."""
    import analysis
    import sweep
    log = ("INFO:root:Training set of size 6912\n"
           "INFO:root:Injected warmup: 0 us\nINFO:root:0: Memory Usage: 900.0, Training Duration: 9.000\n"
           "INFO:root:Injected delay: 0 us\nINFO:root:0: Memory Usage: 900.0, Training Duration: 0.010\n"
           "INFO:root:Injected delay: 25 us\nINFO:root:0: Memory Usage: 900.0, Training Duration: 0.020\n"
           "INFO:root:Injected loss: 0.5 %\nINFO:root:0: Memory Usage: 900.0, Training Duration: 0.030\n"
           "INFO:root:Injected delay: 25 us\nINFO:root:1: Memory Usage: 901.0, Training Duration: 0.025\n")
    parts = sweep.split_network_log(log)
    assert set(parts) == {("delay", "0"), ("delay", "25"), ("loss", "0.5")}
    assert parts[("delay", "25")].count("Training Duration") == 2            # both ranks' segments of the same rule
    assert sweep.tag_fields("loss=0.5") == {"rule_type": "loss", "rule_value": 0.5, "rule": "loss 0.5%"}
    assert sweep.LOSSES_PCT == [0, 0.1, 0.5, 1, 2, 5, 10, 15] and sweep.DELAYS_US[-1] == 400
    cfg = {"trainer": "distributed", "hosts": 8, "slots": 1, "parameters": {"--batch-size": 1440}}
    results = [dict(command=f"c #{k}={v}", stdout="", stderr=text, config=cfg, **sweep.tag_fields(f"{k}={v}"))
               for (k, v), text in parts.items()]
    f = tmp_path / "n.json"
    json.dump({"results": results}, open(f, "w"))
    rows = analysis.load([f])
    t_delay, t_loss = analysis.table(rows, network=True), analysis.table(rows, network=True, rule="loss")
    assert "| distributed | 25.0 | 0.0250 |" in t_delay and "loss" not in t_delay     # max over ranks (Q16)
    assert "| trainer | loss % |" in t_loss and "| distributed | 0.5 | 0.0300 |" in t_loss
    out = analysis.plot(rows, tmp_path / "loss.png", network=True, rule="loss")
    assert Path(out).exists()


def test_deploy_assets_and_launcher_auto(tmp_path):
    """O4-O6 analogues: the compose file parses and describes one all-GPU trainer plus the two-container rendezvous;
    the image bakes no SSH key; `launch -n auto` resolves to one rank per visible GPU (one rank on a CPU machine)."""
    import yaml
    compose = yaml.safe_load(open(ROOT / "deploy" / "docker-compose.yaml"))
    assert set(compose["services"]) == {"trainer", "master", "worker"}
    assert compose["services"]["trainer"]["ipc"] == "host"                    # symmetric memory across the ranks of a node
    for svc in ("master", "worker"):
        assert "B200RNN_COMM=nccl" in compose["services"][svc]["environment"] and "--nnodes 2" in compose["services"][svc]["command"]
    docker = "\n".join(ln for ln in (ROOT / "deploy" / "Dockerfile").read_text().splitlines() if not ln.lstrip().startswith("#"))
    assert "pytorch_distributed_rnn_b200._build" in docker and "id_rsa" not in docker and "ssh" not in docker
    script = tmp_path / "who.py"
    script.write_text("import os; print('rank', os.environ['RANK'], 'of', os.environ['WORLD_SIZE'])")
    cp = subprocess.run([sys.executable, "-m", "pytorch_distributed_rnn_b200.parallel.launch", "-n", "auto", str(script)],
                        capture_output=True, text=True, cwd=ROOT, env={**__import__("os").environ, "CUDA_VISIBLE_DEVICES": ""})
    assert cp.returncode == 0 and cp.stdout.strip() == "rank 0 of 1", cp.stderr


def test_analysis_memory_plot_and_local_baseline(tmp_path):
    """A1 (evaluation/Experiments.ipynb): time AND memory vs ranks per trainer, the local run as a dashed baseline."""
    import analysis
    rows = [dict(trainer="local", ranks=1, batch=1440, seconds=1.0, memory=800.0, samples=6912, rule_type=None, rule_value=None)]
    for n, s_, m in ((1, 1.1, 900.0), (2, 0.6, 950.0), (4, 0.35, 990.0)):
        for rep in (0.0, 0.02):
            rows.append(dict(trainer="distributed", ranks=n, batch=1440, seconds=s_ + rep, memory=m, samples=6912,
                             rule_type=None, rule_value=None))
    t = analysis.table(rows)
    assert "| distributed | 2 | 1440 | 0.6100 |" in t and "| 2 | 0.0141 |" in t          # mean and sd over the repeats
    out_t = Path(analysis.plot(rows, tmp_path / "t.png"))
    out_m = Path(analysis.plot(rows, tmp_path / "m.png", metric="memory"))
    assert out_t.exists() and out_m.exists()
    if out_t.suffix == ".svg":     # the dependency-free renderer: check what it drew
        svg_t, svg_m = out_t.read_text(), out_m.read_text()
        assert "stroke-dasharray" in svg_t and "Training Time (s)" in svg_t and "Memory (MiB)" in svg_m
