#!/usr/bin/env python
"""Gradient all-reduce latency / bandwidth sweep, 4 KB - 256 MB, in-kernel NVLink vs NCCL.

This is the analogue of the reference's network experiment (fabfile.py:130-191,
evaluation/Experiments_network.ipynb): there, `tc netem` delay/loss was injected on 12 Pis and
the epoch time of DDP vs Horovod was recorded; here the interconnect is NVLink 5 / NVSwitch, so
the experiment becomes (a) the latency/bandwidth curve of the gradient all-reduce itself for each
algorithm (one-shot / two-shot / NCCL) and (b) the same curve with an injected delay in front of
every flag publication (``--delay-us``), which shows how each algorithm's critical path multiplies
a per-message latency (one-shot: x1, two-shot: x2 — cf. DDP's ~25 serialized latencies per step and
Horovod's extra negotiation rounds).

Launch (N ranks):  torchrun --nproc-per-node N bench/allreduce_sweep.py [--out results.json]
Timing: CUDA events on the launching stream around 8 back-to-back calls, per size 20 such
measurements after 5 warm-ups, median, max over ranks.  Roofline per size: bytes that must cross NVLink per
rank over the link bandwidth, reported against BOTH 770 GB/s (measured peer-copy bandwidth,
B200_PROFILING.md) and 900 GB/s (nominal NVLink 5 per direction); one-shot sends (W-1)*S (the LL form
2*(W-1)*S: every word carries its sequence number), two-shot 2*(W-1)/W*S.

``--write-table`` turns the measured crossovers into ``profiles/allreduce_table.json`` (merged per world
size), which ``SymmComm`` loads to pick the variant per bucket size; ``--md`` writes the table as markdown.
"""
from __future__ import annotations

import argparse
import json
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from pytorch_distributed_rnn_b200 import FusedSGD, _native  # noqa: E402
from pytorch_distributed_rnn_b200.parallel import process_group as pg  # noqa: E402
from pytorch_distributed_rnn_b200.parallel.symm import SymmComm  # noqa: E402

NVLINK_GBS = 770.0        # measured peer copy
NVLINK_NOMINAL_GBS = 900.0
ALGOS = ((4, "ll"), (1, "oneshot"), (2, "twoshot"), (3, "nvls"))


def crossovers(rows):
    """Measured crossovers (bytes) between the variants for one world size, rows sorted by size, delay 0."""
    rows = sorted((r for r in rows if r["delay_us"] == 0 and not r.get("loss_pct")), key=lambda r: r["bytes"])
    ll_max, one_max, nvls_min = 0, 0, None
    for r in rows:
        if "oneshot_us" in r and "twoshot_us" in r and r["oneshot_us"] <= r["twoshot_us"]:
            one_max = max(one_max, r["bytes"])
    # the LL range must be contiguous from the smallest size
    for r in rows:
        ours = {k: r[f"{k}_us"] for _, k in ALGOS if f"{k}_us" in r}
        if "ll" in ours and min(ours, key=ours.get) == "ll":
            ll_max = r["bytes"]
        else:
            break
    for r in reversed(rows):   # NVLS from the smallest size on which it wins everywhere above
        if "nvls_us" in r and "twoshot_us" in r and r["nvls_us"] <= r["twoshot_us"]:
            nvls_min = r["bytes"]
        else:
            break
    return {"ll_max_bytes": ll_max, "oneshot_max_bytes": one_max,
            "nvls_min_bytes": nvls_min if nvls_min is not None else 1 << 62}


def write_md(path, results, world):
    lines = [f"# All-reduce sweep, W = {world} (fused reduce + all-reduce + optimizer epilogue; us, median, max over ranks)", "",
             "| bytes | rows | LL | one-shot | two-shot | NVLS | NCCL all_reduce | best own / NCCL | best-own frac of roofline @770 | @900 |",
             "|---|---|---|---|---|---|---|---|---|---|"]
    for r in results:
        if r["delay_us"] != 0 or r.get("loss_pct"):
            continue
        ours = {k: r[f"{k}_us"] for _, k in ALGOS if f"{k}_us" in r}
        best = min(ours, key=ours.get)
        f = lambda k: f"{r[k + '_us']:.1f}" if k + "_us" in r else "-"
        roof = r["roofline_twoshot_us"] if best in ("twoshot", "nvls") else (2 if best == "ll" else 1) * r["roofline_oneshot_us"]
        lines.append(f"| {r['bytes']} | {r['rows']} | {f('ll')} | {f('oneshot')} | {f('twoshot')} | {f('nvls')} | {f('nccl')} | "
                     f"{best} {ours[best] / r['nccl_us']:.2f}x | {roof / ours[best]:.3f} | {roof * NVLINK_GBS / NVLINK_NOMINAL_GBS / ours[best]:.3f} |")
    delayed = [r for r in results if r["delay_us"] != 0]
    if delayed:
        lines += ["", "## Injected delay before every publication (the tc-netem analogue)", "",
                  "| delay us | bytes | LL | one-shot | two-shot | NVLS |", "|---|---|---|---|---|---|"]
        for r in delayed:
            f = lambda k: f"{r[k + '_us']:.1f}" if k + "_us" in r else "-"
            lines.append(f"| {r['delay_us']} | {r['bytes']} | {f('ll')} | {f('oneshot')} | {f('twoshot')} | {f('nvls')} |")
    lossy = [r for r in results if r.get("loss_pct")]
    if lossy:
        lines += ["", "## Injected message loss (a dropped publication pays a 200 us retransmission timeout)", "",
                  "| loss % | bytes | LL | one-shot | two-shot | NVLS |", "|---|---|---|---|---|---|"]
        for r in lossy:
            f = lambda k: f"{r[k + '_us']:.1f}" if k + "_us" in r else "-"
            lines.append(f"| {r['loss_pct']} | {r['bytes']} | {f('ll')} | {f('oneshot')} | {f('twoshot')} | {f('nvls')} |")
    Path(path).write_text("\n".join(lines) + "\n")


REPS = 8   # calls per event pair: the GPU queue stays full, so host launch cost (~20 us of Python) is not timed


def timed(fn, iters, warm, dev):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize(dev)
    dist.barrier()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(REPS):
            fn()
        b.record()
        torch.cuda.synchronize(dev)
        ts.append(a.elapsed_time(b) * 1e3 / REPS)
    t = torch.tensor([sorted(ts)[len(ts) // 2]], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--min-bytes", type=int, default=4 << 10)
    ap.add_argument("--max-bytes", type=int, default=256 << 20)
    ap.add_argument("--delay-us", type=str, default="0", help="comma list of injected delays (us)")
    ap.add_argument("--loss-pct", type=str, default="", help="comma list of injected message-loss percentages (run after the delays)")
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--rows", type=int, default=1, help="local gradient rows to reduce inside the call (the fused step has ~90-144)")
    ap.add_argument("--out", default=None)
    ap.add_argument("--md", default=None)
    ap.add_argument("--write-table", default=None, help="merge the measured crossovers into this JSON (profiles/allreduce_table.json)")
    ap.add_argument("--extra-bytes", default="56600", help="comma list of extra sizes (the reference model's gradient is 56 600 B)")
    a = ap.parse_args()
    import os
    os.environ.setdefault("B200RNN_LL_FLOATS", str(1 << 20))   # let the sweep try LL up to 4 MB
    info = pg.init_distributed()
    rank, world, dev = info.rank, info.world_size, info.device
    max_floats = a.max_bytes // 4
    comm = SymmComm(info, inbox_floats=max_floats, heap_bytes=2 * a.max_bytes + (64 << 20), max_blocks=1 << 18)
    lib = _native.lib()
    results = []
    sizes = []
    b = a.min_bytes
    while b <= a.max_bytes:
        sizes.append(b)
        b *= 4
    sizes = sorted(set(sizes + [int(x) for x in a.extra_bytes.split(",") if x and a.min_bytes <= int(x) <= a.max_bytes]))
    rules = [(int(x), 0.0) for x in a.delay_us.split(",")] + [(0, float(x)) for x in a.loss_pct.split(",") if x]
    for delay, loss in rules:
        comm.set_delay_ns(delay * 1000)
        comm.set_loss(loss)
        for nbytes in sizes:
            n = nbytes // 4
            # the vector lives in the symmetric heap (needed by two-shot's parameter all-gather)
            lin = torch.nn.Linear(n - 1, 1, bias=True, device=dev) if n > 1 else None
            arena = comm.make_arena(lin)
            opt = FusedSGD(arena, lr=0.0)
            rows = torch.ones(a.rows, n + 2, device=dev)
            slot = torch.zeros(2, device=dev)
            row = {"bytes": nbytes, "world": world, "delay_us": delay, "loss_pct": loss, "rows": a.rows}
            for algo, name in ALGOS:
                if algo == 3 and comm.mc_base is None:
                    continue
                if algo == 1 and (n + 1023) // 1024 > comm.ctx.max_blocks:
                    continue
                if algo == 4 and (n + 3) // 4 * 4 > comm.ctx.ll_floats:
                    continue
                comm.algo = algo

                def fn():
                    opt.step_count += 1
                    comm.allreduce_update(rows, a.rows, arena, opt, opt.desc(grad_scale=1.0 / world), slot)
                row[f"{name}_us"] = timed(fn, a.iters, 5, dev)
            comm.algo = 0
            if delay == 0 and loss == 0:
                t = torch.ones(n, device=dev)
                row["nccl_us"] = timed(lambda: dist.all_reduce(t), a.iters, 5, dev)
            per_rank_one = (world - 1) * nbytes
            per_rank_two = 2 * (world - 1) / world * nbytes
            row["roofline_oneshot_us"] = per_rank_one / (NVLINK_GBS * 1e3)
            row["roofline_twoshot_us"] = per_rank_two / (NVLINK_GBS * 1e3)
            for name in ("ll", "oneshot", "twoshot", "nvls", "nccl"):
                if f"{name}_us" in row:
                    roof = {"oneshot": row["roofline_oneshot_us"], "ll": 2 * row["roofline_oneshot_us"]}.get(name, row["roofline_twoshot_us"])
                    row[f"{name}_frac_of_nominal_900"] = roof * NVLINK_GBS / NVLINK_NOMINAL_GBS / row[f"{name}_us"] if row[f"{name}_us"] > 0 else None
                    row[f"{name}_frac_of_nvlink_roofline"] = roof / row[f"{name}_us"] if row[f"{name}_us"] > 0 else None
                    row[f"{name}_busbw_GBs"] = 2 * (world - 1) / world * nbytes / row[f"{name}_us"] / 1e3
            comm.check()
            results.append(row)
            if rank == 0:
                print(json.dumps(row), flush=True)
            # free the heap slot of this size (bump allocator: rewind, sizes only grow)
            comm._heap_top = arena.heap_off
    if rank == 0 and a.out:
        Path(a.out).parent.mkdir(parents=True, exist_ok=True)
        json.dump({"results": results, "nvlink_GBs": NVLINK_GBS, "nvlink_nominal_GBs": NVLINK_NOMINAL_GBS,
                   "note": "fused = reduce the local rows + all-reduce + SGD(lr=0) epilogue, in one kernel"}, open(a.out, "w"), indent=1)
    if rank == 0 and a.md:
        write_md(a.md, results, world)
    if rank == 0 and a.write_table:
        p = Path(a.write_table)
        doc = json.loads(p.read_text()) if p.exists() else {"worlds": {}}
        doc.setdefault("worlds", {})[str(world)] = dict(crossovers(results), measured_rows=a.rows)
        doc["note"] = ("crossovers (vector bytes) between the in-kernel all-reduce variants, measured by bench/allreduce_sweep.py; "
                       "loaded by pytorch_distributed_rnn_b200.parallel.symm.load_algo_table")
        p.parent.mkdir(parents=True, exist_ok=True)
        p.write_text(json.dumps(doc, indent=1) + "\n")
    dist.barrier()
    comm.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
