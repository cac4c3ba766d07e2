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


def table(rows, network=False, rule="delay"):
    if network:
        rows = [r for r in rows if (r.get("rule_type") or "delay") == rule]
    key = (lambda r: (r["trainer"], r["rule_value"])) if network else (lambda r: (r["trainer"], r["ranks"], r["batch"]))
    groups = defaultdict(list)
    for r in rows:
        groups[key(r)].append(r)
    out = []
    head = (f"| trainer | {'loss %' if rule == 'loss' else 'delay us'} | s | samples/s | n | sd s |" if network
            else "| trainer | ranks | batch | s | samples/s | MiB | n | sd s |")
    out.append(head)
    out.append("|" + "---|" * (head.count("|") - 1))
    for k in sorted(groups, key=lambda x: tuple(str(v) if v is None else v for v in x)):
        g = groups[k]
        s = statistics.mean(x["seconds"] for x in g)
        sps = g[0]["samples"] / s
        sd = f"{statistics.stdev(x['seconds'] for x in g):.4f}" if len(g) > 1 else "-"   # the notebooks draw a 68 % band
        if network:
            out.append(f"| {k[0]} | {k[1]} | {s:.4f} | {sps:.0f} | {len(g)} | {sd} |")
        else:
            out.append(f"| {k[0]} | {k[1]} | {k[2]} | {s:.4f} | {sps:.0f} | {statistics.mean(x['memory'] for x in g):.0f} | {len(g)} | {sd} |")
    return "\n".join(out)


def _series(rows, network, metric="seconds"):
    xs = "rule_value" if network else "ranks"
    out = {}
    for tr in sorted({r["trainer"] for r in rows}):
        pts = defaultdict(list)
        for r in rows:
            if r["trainer"] == tr and r[xs] is not None:
                pts[r[xs]].append(r[metric])
        if pts:
            k = sorted(pts)
            out[tr] = (k, [statistics.mean(pts[x]) for x in k],
                       [statistics.stdev(pts[x]) if len(pts[x]) > 1 else 0.0 for x in k])
    return out


def plot(rows, path, network=False, rule="delay", metric="seconds"):
    """Training time (or peak memory, ``metric="memory"``) vs ranks (or injected delay / loss) per trainer, the `local`
    trainer as a dashed horizontal baseline like in Experiments.ipynb.  matplotlib (Agg) when it is installed, else a
    dependency-free SVG with the same content (``path`` then gets the .svg suffix); returns the file written."""
    if network:
        rows = [r for r in rows if (r.get("rule_type") or "delay") == rule]
    series = _series(rows, network, metric)
    ylabel = "Memory (MiB)" if metric == "memory" else "Training Time (s)"
    xlabel = ("injected loss (%)" if rule == "loss" else "injected delay (us)") if network else "ranks (GPUs)"
    try:
        import matplotlib
        matplotlib.use("Agg")
        import matplotlib.pyplot as plt
    except ImportError:
        return _plot_svg(series, Path(path).with_suffix(".svg"), xlabel, ylabel)
    fig, ax = plt.subplots(figsize=(6, 4))
    for tr, (k, v, sd) in series.items():
        if len(k) == 1 and len(series) > 1:      # the single-rank baseline
            ax.axhline(v[0], linestyle="--", color="gray", label=tr)
            continue
        ax.errorbar(k, v, yerr=sd if any(sd) else None, marker="o", capsize=3, label=tr)   # +-1 sd = the notebooks' 68 % band
    ax.set_xlabel(xlabel)
    ax.set_ylabel(ylabel)
    ax.legend()
    fig.tight_layout()
    fig.savefig(path, dpi=120)
    return path


def _plot_svg(series, path, xlabel, ylabel="Training Time (s)", w=640, h=420):
    ml, mr, mt, mb = 70, 20, 20, 50
    xs = sorted({x for k, _, _ in series.values() for x in k})
    ys = [y + e for _, v, sd in series.values() for y, e in zip(v, sd)]
    if not xs or not ys:
        return None
    x0, x1, y1 = min(xs), max(xs), max(ys) * 1.08
    sx = lambda x: ml + (w - ml - mr) * ((x - x0) / (x1 - x0) if x1 > x0 else 0.5)
    sy = lambda y: h - mb - (h - mt - mb) * (y / y1 if y1 > 0 else 0.0)
    colors = ["#1f77b4", "#ff7f0e", "#2ca02c", "#d62728", "#9467bd"]
    o = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{w}" height="{h}" font-family="sans-serif" font-size="12">',
         f'<rect width="{w}" height="{h}" fill="white"/>',
         f'<line x1="{ml}" y1="{h - mb}" x2="{w - mr}" y2="{h - mb}" stroke="black"/>',
         f'<line x1="{ml}" y1="{mt}" x2="{ml}" y2="{h - mb}" stroke="black"/>']
    for x in xs:
        o.append(f'<text x="{sx(x):.1f}" y="{h - mb + 16}" text-anchor="middle">{x:g}</text>')
    for i in range(6):
        y = y1 * i / 5
        o.append(f'<text x="{ml - 6}" y="{sy(y) + 4:.1f}" text-anchor="end">{y:.4g}</text>')
        o.append(f'<line x1="{ml}" y1="{sy(y):.1f}" x2="{w - mr}" y2="{sy(y):.1f}" stroke="#dddddd"/>')
    o.append(f'<text x="{(ml + w - mr) / 2}" y="{h - 10}" text-anchor="middle">{xlabel}</text>')
    o.append(f'<text x="16" y="{(mt + h - mb) / 2}" text-anchor="middle" transform="rotate(-90 16 {(mt + h - mb) / 2})">{ylabel}</text>')
    for i, (tr, (k, v, sd)) in enumerate(series.items()):
        c = colors[i % len(colors)]
        pts = " ".join(f"{sx(x):.1f},{sy(y):.1f}" for x, y in zip(k, v))
        if len(k) == 1 and len(series) > 1:      # the single-rank baseline: dashed, across the plot
            o.append(f'<line x1="{ml}" y1="{sy(v[0]):.1f}" x2="{w - mr}" y2="{sy(v[0]):.1f}" stroke="{c}" stroke-dasharray="6 4"/>')
        o.append(f'<polyline points="{pts}" fill="none" stroke="{c}" stroke-width="2"/>')
        for x, y, e in zip(k, v, sd):
            o.append(f'<circle cx="{sx(x):.1f}" cy="{sy(y):.1f}" r="3.5" fill="{c}"/>')
            if e > 0:   # +-1 standard deviation over the repeats
                o.append(f'<line x1="{sx(x):.1f}" y1="{sy(y - e):.1f}" x2="{sx(x):.1f}" y2="{sy(y + e):.1f}" stroke="{c}"/>')
        o.append(f'<rect x="{w - mr - 130}" y="{mt + 18 * i}" width="12" height="12" fill="{c}"/>')
        o.append(f'<text x="{w - mr - 112}" y="{mt + 18 * i + 11}">{tr}</text>')
    o.append("</svg>")
    Path(path).write_text("\n".join(o) + "\n")
    return path


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("files", nargs="+")
    ap.add_argument("--plot", default=None)
    ap.add_argument("--network", action="store_true")
    ap.add_argument("--metric", default="seconds", choices=["seconds", "memory"], help="what --plot draws (Experiments.ipynb plots both)")
    ap.add_argument("--rule", default="delay", choices=["delay", "loss"],
                    help="--network: which rule type to tabulate (Experiments_network.ipynb plots both)")
    a = ap.parse_args()
    rows = load(a.files)
    print(table(rows, a.network, a.rule))
    if a.plot:
        plot(rows, a.plot, a.network, a.rule, a.metric)


if __name__ == "__main__":
    main()
