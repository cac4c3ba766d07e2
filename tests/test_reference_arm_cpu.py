"""The reference arm's plumbing that does not need a GPU: installing the unmodified reference on demand, and the
network experiment's log splitting."""
import importlib
import json
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parent.parent


def test_ensure_reference_populates_ref_dir_from_a_source_tree(tmp_path, monkeypatch):
    """A fresh checkout has no baseline/_ref: ensure_reference() must create it (here from a stand-in source tree and
    through the plain-copy fallback, pip being pointed at nothing installable)."""
    src = tmp_path / "reference"
    for rel, body in (("src/motion/trainer/base.py", "MARK = 'base'\n"), ("src/motion/model.py", "MARK = 'model'\n"),
                      ("src/example/example_ddp.py", "MARK = 'ddp'\n")):
        f = src / rel
        f.parent.mkdir(parents=True, exist_ok=True)
        f.write_text(body)
    import baseline
    monkeypatch.setattr(baseline, "REF_SRC", src)
    monkeypatch.setattr(baseline, "REF_ROOT", tmp_path / "_ref")
    monkeypatch.setattr(baseline, "_MARKER", tmp_path / "_ref" / "motion" / "trainer" / "base.py")
    monkeypatch.setattr(baseline, "HERE", tmp_path)          # no install script here -> pip path fails -> copy fallback
    how = baseline.ensure_reference()
    assert "copy" in how
    assert (tmp_path / "_ref" / "motion" / "model.py").read_text() == "MARK = 'model'\n"
    assert (tmp_path / "_ref" / "example" / "example_ddp.py").exists()
    assert baseline.ensure_reference() == "already installed"


def test_real_reference_is_installed_byte_for_byte_when_available():
    ref = Path("/root/reference/src/motion")
    if not ref.exists():
        return
    import baseline
    baseline.ensure_reference()
    for f in ("model.py", "trainer/base.py", "trainer/distributed.py", "dataset.py"):
        assert (baseline.REF_ROOT / "motion" / f).read_bytes() == (ref / f).read_bytes(), f


def test_network_sweep_splits_one_log_into_per_delay_entries(tmp_path, monkeypatch):
    sys.path.insert(0, str(ROOT / "bench"))
    sweep = importlib.import_module("sweep")
    log = ""
    for rank in (0, 1):
        for d, dur in ((0, 0.002), (50, 0.0025), (400, 0.004)):
            log += f"INFO:root:Injected delay: {d} us\nINFO:root:{rank}: Memory Usage: 100.0, Training Duration: {dur + rank * 1e-4}\n"

    class CP:
        returncode, stdout, stderr = 0, "", log
    monkeypatch.setattr(sweep.subprocess, "run", lambda *a, **k: CP())

    class A:
        dataset_path, results, delays, losses = tmp_path / "ds", str(tmp_path / "net.json"), "0,50,400", ""
    sweep.run_network(A(), 2)
    res = json.load(open(A.results))["results"]
    assert len(res) == 6                                  # 3 delays x 2 trainers
    by = {(r["trainer"], r["rule_value"]): r for r in res}
    assert by[("distributed", 400.0)]["rule"] == "delay 400us" and by[("horovod", 50.0)]["rule_type"] == "delay"
    assert by[("distributed", 50.0)]["stderr"].count("Training Duration") == 2      # both ranks' lines of that delay
    sys.path.insert(0, str(ROOT))
    analysis = importlib.import_module("analysis")
    rows = analysis.load([A.results])
    got = {(r["trainer"], r["rule_value"]): r["seconds"] for r in rows}
    assert abs(got[("distributed", 400.0)] - 0.0041) < 1e-9   # max over ranks
    sweep.run_network(A(), 2)                             # resume: nothing is run or recorded twice
    assert len(json.load(open(A.results))["results"]) == 6


def test_bench_prints_exactly_one_json_line_without_a_gpu():
    """bench.py's stdout contract (one JSON line, exit 0) also holds where it cannot measure: without a CUDA device both
    arms say so instead of crashing or printing library chatter."""
    import subprocess
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is visible: covered by the real bench runs")
    for extra in ([], ["--impl", "reference"], ["--config", "scaled"]):
        cp = subprocess.run([sys.executable, str(ROOT / "bench.py"), "--steps", "1", "--warmup", "3", *extra],
                            capture_output=True, text=True, timeout=300, cwd=ROOT)
        assert cp.returncode == 0, cp.stderr[-2000:]
        lines = [l for l in cp.stdout.splitlines() if l.strip()]
        assert len(lines) == 1, cp.stdout
        d = json.loads(lines[0])
        assert d["impl"] == ("reference" if "reference" in extra else "own") and "CUDA" in d["unavailable"]
