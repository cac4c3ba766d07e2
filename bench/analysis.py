#!/usr/bin/env python
"""Tables / plots from sweep results — successor of evaluation/Experiments*.ipynb.

Parses the reference's result schema (and therefore also the reference's own result files, e.g.
/root/reference/evaluation/results_202007141730.json) with the notebooks' regex on stderr
(``(\\d+): Memory Usage: (\\d+\\.\\d+), Training Duration: (\\d+\\.\\d+)``, Experiments.ipynb:49),
but takes the MAX over ranks instead of the first match (quirk Q16) and de-duplicates commands
(Q17).  Prints a markdown table (trainer x ranks x batch size -> seconds, samples/s, memory) and,
with ``--plot out.png``, draws time vs ranks per trainer (matplotlib Agg, file output only).

    python bench/analysis.py results.json [more.json ...] [--plot scaling.png] [--network]
"""
from __future__ import annotations

import argparse
import json
import re
import statistics
from collections import defaultdict

import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200.trainer.formatter import parse_log  # noqa: E402  (the log-line contract's parsers)

SIZE_RE = re.compile(r"Training set of size (\d+)")


def load(files):
    seen, rows = set(), []
    for f in files:
        with open(f) as fh:
            for r in json.load(fh)["results"]:
                if r["command"] in seen:
                    continue
                seen.add(r["command"])
                perf = [rec for rec in parse_log(r.get("stderr", "")) if rec["kind"] == "performance"]
                if not perf:
                    continue
                m = SIZE_RE.search(r.get("stderr", ""))
                cfg = r.get("config", {})
                params = cfg.get("parameters", {})
                rows.append({
                    "trainer": cfg.get("trainer", r.get("trainer")), "ranks": cfg.get("hosts", 0) * cfg.get("slots", 1) or None,
                    "batch": params.get("--batch-size", 1440), "seconds": max(p["duration"] for p in perf),
                    "memory": max(p["memory"] for p in perf), "samples": int(m.group(1)) if m else 6912,
                    "rule_type": r.get("rule_type"), "rule_value": r.get("rule_value")})
    return rows


def table(rows, network=False):
    key = (lambda r: (r["trainer"], r["rule_value"])) if network else (lambda r: (r["trainer"], r["ranks"], r["batch"]))
    groups = defaultdict(list)
    for r in rows:
        groups[key(r)].append(r)
    out = []
    head = "| trainer | delay us | s | samples/s | n |" if network else "| trainer | ranks | batch | s | samples/s | MiB | n |"
    out.append(head)
    out.append("|" + "---|" * (head.count("|") - 1))
    for k in sorted(groups, key=lambda x: tuple(str(v) if v is None else v for v in x)):
        g = groups[k]
        s = statistics.mean(x["seconds"] for x in g)
        sps = g[0]["samples"] / s
        if network:
            out.append(f"| {k[0]} | {k[1]} | {s:.4f} | {sps:.0f} | {len(g)} |")
        else:
            out.append(f"| {k[0]} | {k[1]} | {k[2]} | {s:.4f} | {sps:.0f} | {statistics.mean(x['memory'] for x in g):.0f} | {len(g)} |")
    return "\n".join(out)


def plot(rows, path, network=False):
    import matplotlib
    matplotlib.use("Agg")
    import matplotlib.pyplot as plt
    fig, ax = plt.subplots(figsize=(6, 4))
    xs = "rule_value" if network else "ranks"
    for tr in sorted({r["trainer"] for r in rows}):
        pts = defaultdict(list)
        for r in rows:
            if r["trainer"] == tr and r[xs] is not None:
                pts[r[xs]].append(r["seconds"])
        if pts:
            k = sorted(pts)
            ax.plot(k, [statistics.mean(pts[x]) for x in k], marker="o", label=tr)
    ax.set_xlabel("injected delay (us)" if network else "ranks (GPUs)")
    ax.set_ylabel("Training Time (s)")
    ax.legend()
    fig.tight_layout()
    fig.savefig(path, dpi=120)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("files", nargs="+")
    ap.add_argument("--plot", default=None)
    ap.add_argument("--network", action="store_true")
    a = ap.parse_args()
    rows = load(a.files)
    print(table(rows, a.network))
    if a.plot:
        plot(rows, a.plot, a.network)


if __name__ == "__main__":
    main()
