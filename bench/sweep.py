#!/usr/bin/env python
"""Experiment grid runner — the fabfile analogue (fabfile.py:48-66,194-290).

Grid: batch size {480, 960, 1440} x ranks {1, 2, 4, 8} (one rank per GPU; the reference used
1/2/4/8/12 Pi hosts) x trainer {local (1 rank only), distributed, horovod}, every run with
``--epochs 1 --seed 123456789 --no-validation`` like TRAIN_RUNS.  Behaviour kept from
``run_training``: configurations are shuffled, a command already present in the results file
is skipped (experiment-level resume), a failed run is recorded rather than aborting the sweep,
and the file is rewritten after every run in the reference's schema
``{"results": [{command, stdout, stderr, config{trainer,hosts,slots,parameters}}]}`` so the
analysis (bench/analysis.py, the notebooks' successor) parses both.  Fixes: the results file is
created if missing (quirk Q15).  ``--network`` runs the delay-injection variant (O2): the
``distributed`` and ``horovod`` trainers at the largest rank count for each injected delay.

    python bench/sweep.py --dataset-path /tmp/motion --results results.json [--ranks 1,2,4,8] [--repeats 2]

``--pack G`` (G = GPUs of the box) runs configurations that need fewer than G GPUs side by side on disjoint
GPU sets (CUDA_VISIBLE_DEVICES), largest world first — the reference ran its grid on one cluster serially;
here a run is a few seconds of start-up around a few milliseconds of training, so packing the box is what keeps
the experiment affordable.  The recorded duration is each run's own in-process measurement either way.
"""
from __future__ import annotations

import argparse
import json
import os
import random
import subprocess
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
BATCH_SIZES = [480, 960, 1440]
TRAINERS = ["local", "distributed", "horovod"]
DELAYS_US = [0, 1, 2, 5, 10, 25, 50, 100, 200, 300, 400]   # the reference's delay list, ms -> us (NVLink scale)
LOSSES_PCT = [0, 0.1, 0.5, 1, 2, 5, 10, 15]                 # the reference's loss list (fabfile.py:149-151 and results)


def train_runs(ranks, repeats):
    return [{"trainer": tr, "hosts": n, "slots": 1, "repeat": rep,
             "parameters": {"--batch-size": bs, "--epochs": 1, "--seed": 123456789, "--no-validation": ""}}
            for bs in BATCH_SIZES for n in ranks for tr in TRAINERS for rep in range(repeats)
            if tr != "local" or n == 1]


def get_command(run, dataset_path, extra=()):
    params = []
    for k, v in run["parameters"].items():
        params += [k] if v == "" else [k, str(v)]
    base = ["--dataset-path", str(dataset_path), "--synthetic", "--deferred-logging", *params, *extra, run["trainer"]]
    n = run["hosts"] * run["slots"]
    if run["trainer"] == "local":
        return [sys.executable, "-m", "pytorch_distributed_rnn_b200", *base]
    return [sys.executable, "-m", "pytorch_distributed_rnn_b200.parallel.launch", "-n", str(n),
            "-m", "pytorch_distributed_rnn_b200", *base]


def load_results(path):
    if not Path(path).exists():
        return []
    with open(path) as f:
        return json.load(f)["results"]


def run_training(runs, dataset_path, results_file, env=None, tag=None):
    runs = list(runs)
    random.shuffle(runs)
    for i, run in enumerate(runs):
        cmd = get_command(run, dataset_path)
        command = " ".join(cmd) + (f"  #{tag}" if tag else "") + f"  #repeat{run.get('repeat', 0)}"
        results = load_results(results_file)
        if any(r["command"] == command for r in results):
            print(f"[{i + 1}/{len(runs)}] skip (already recorded): {command}")
            continue
        print(f"[{i + 1}/{len(runs)}] {command}", flush=True)
        e = dict(os.environ, PYTHONPATH=str(ROOT))
        e.update(env or {})
        cp = subprocess.run(cmd, capture_output=True, text=True, env=e, cwd=str(ROOT))  # warn=True semantics
        entry = {"command": command, "stdout": cp.stdout, "stderr": cp.stderr, "returncode": cp.returncode,
                 "config": {k: run[k] for k in ("trainer", "hosts", "slots", "parameters")}}
        if tag:
            entry.update(tag_fields(tag))
        results.append(entry)
        with open(results_file, "w") as f:
            json.dump({"results": results}, f)


def run_packed(runs, dataset_path, results_file, n_gpus, env=None, tag=None):
    """Same bookkeeping as run_training, with runs of n ranks started n_gpus // n at a time on disjoint GPUs."""
    import tempfile
    import time
    by_n = {}
    for run in runs:
        by_n.setdefault(run["hosts"] * run["slots"], []).append(run)
    for n in sorted(by_n, reverse=True):
        pending = list(by_n[n])
        random.shuffle(pending)
        free = [list(range(i * n, (i + 1) * n)) for i in range(max(1, n_gpus // n))]
        active = []
        while pending or active:
            while pending and free:
                run = pending.pop()
                cmd = get_command(run, dataset_path)
                command = " ".join(cmd) + (f"  #{tag}" if tag else "") + f"  #repeat{run.get('repeat', 0)}"
                if any(r["command"] == command for r in load_results(results_file)):
                    print(f"skip (already recorded): {command}")
                    continue
                gpus = free.pop()
                e = dict(os.environ, PYTHONPATH=str(ROOT), CUDA_VISIBLE_DEVICES=",".join(map(str, gpus)))
                e.update(env or {})
                fo, fe = tempfile.TemporaryFile("w+"), tempfile.TemporaryFile("w+")
                print(f"[n={n} gpus={gpus}] {command}", flush=True)
                active.append((run, command, gpus, subprocess.Popen(cmd, stdout=fo, stderr=fe, env=e, cwd=str(ROOT)), fo, fe))
            for item in list(active):
                run, command, gpus, proc, fo, fe = item
                if proc.poll() is None:
                    continue
                fo.seek(0); fe.seek(0)
                entry = {"command": command, "stdout": fo.read(), "stderr": fe.read(), "returncode": proc.returncode,
                         "config": {k: run[k] for k in ("trainer", "hosts", "slots", "parameters")}, "gpus": gpus}
                if tag:
                    entry.update(tag_fields(tag))
                results = load_results(results_file)
                results.append(entry)
                with open(results_file, "w") as f:
                    json.dump({"results": results}, f)
                active.remove(item)
                free.append(gpus)
            time.sleep(0.1)


def tag_fields(tag):
    kind, value = tag.split("=")
    return {"rule_type": kind, "rule_value": float(value), "rule": f"{kind} {value}{'%' if kind == 'loss' else 'us'}"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--dataset-path", default="/tmp/b200rnn_motion")
    ap.add_argument("--results", default="results.json")
    ap.add_argument("--ranks", default="1,2,4,8")
    ap.add_argument("--repeats", type=int, default=2)
    ap.add_argument("--network", action="store_true", help="delay-injection experiment instead of the scaling grid")
    ap.add_argument("--pack", type=int, default=0, help="GPUs of the box: run smaller worlds side by side on disjoint GPUs")
    ap.add_argument("--delays", default=",".join(map(str, DELAYS_US)), help="--network: injected delays in us")
    ap.add_argument("--losses", default=",".join(map(str, LOSSES_PCT)), help="--network: injected message loss in percent")
    a = ap.parse_args()
    ranks = [int(x) for x in a.ranks.split(",")]
    # one dataset for the whole experiment, written before anything runs (runs may start side by side)
    sys.path.insert(0, str(ROOT))
    from pytorch_distributed_rnn_b200.data import synthetic
    if not synthetic.is_complete(a.dataset_path):
        synthetic.write_dataset(a.dataset_path, seed=123456789)
    if not a.network:
        if a.pack:
            run_packed(train_runs(ranks, a.repeats), a.dataset_path, a.results, a.pack)
        else:
            run_training(train_runs(ranks, a.repeats), a.dataset_path, a.results)
        return
    n = max(ranks)
    run_network(a, n)


def split_network_log(log):
    """{(rule_type, value string): log text} from a `--delay-sweep-us` / `--loss-sweep-pct` job's stderr: every rank's
    log contributes one segment per rule, behind its `Injected delay: N us` / `Injected loss: P %` marker."""
    import re
    parts = re.split(r"^.*Injected (delay|loss|warmup): ([0-9.]+) (?:us|%).*$", log, flags=re.M)
    by_rule = {}
    for kind, value, text in zip(parts[1::3], parts[2::3], parts[3::3]):
        if kind != "warmup":      # the sweep's untimed first run
            by_rule[(kind, value)] = by_rule.get((kind, value), "") + text
    return by_rule


def run_network(a, n):
    """O2: the delay / loss experiment.  One job per trainer sweeps all rules in-process (`--delay-sweep-us`,
    `--loss-sweep-pct`: start-up costs seconds, an epoch milliseconds); its log is split at the rule markers into one
    result entry per (trainer, rule), in the reference's network-results schema."""
    for tr in ("distributed", "horovod"):
        params = {"--batch-size": 1440, "--epochs": 1, "--seed": 123456789, "--no-validation": ""}
        if a.delays:
            params["--delay-sweep-us"] = a.delays
        if a.losses:
            params["--loss-sweep-pct"] = a.losses
        run = {"trainer": tr, "hosts": n, "slots": 1, "repeat": 0, "parameters": params}
        cmd = get_command(run, a.dataset_path)
        command = " ".join(cmd)
        results = load_results(a.results)
        if any(r.get("sweep_command") == command for r in results):
            print(f"skip (already recorded): {command}")
            continue
        print(command, flush=True)
        cp = subprocess.run(cmd, capture_output=True, text=True, env=dict(os.environ, PYTHONPATH=str(ROOT)), cwd=str(ROOT))
        by_rule = split_network_log(cp.stderr)
        for (kind, value), log in by_rule.items():
            entry = {"command": f"{command}  #{kind}={value}", "sweep_command": command, "trainer": tr, "stdout": "", "stderr": log,
                     "returncode": cp.returncode, "config": {k: run[k] for k in ("trainer", "hosts", "slots", "parameters")}}
            entry.update(tag_fields(f"{kind}={value}"))
            results.append(entry)
        if not by_rule:   # the job failed before the first marker: record it as the reference records failed runs
            results.append({"command": command, "sweep_command": command, "trainer": tr, "stdout": cp.stdout, "stderr": cp.stderr,
                            "returncode": cp.returncode, "config": {k: run[k] for k in ("trainer", "hosts", "slots", "parameters")}})
        with open(a.results, "w") as f:
            json.dump({"results": results}, f)


if __name__ == "__main__":
    main()
