#!/usr/bin/env python
"""Headline benchmark: motion-LSTM data-parallel training throughput (samples/s).

Config (BASELINE.json / BASELINE.md): the reference's motion classifier — LSTM 9->32, 2 layers,
T=128, Linear 32->6, CrossEntropy, Adam lr 0.0025, fp32 — at GLOBAL batch 1440 (per-rank batch
1440 // N, the reference's rule, src/motion/trainer/distributed.py:48-49 => strong scaling), on
synthetic data of the reference's shape (6 912 x [128, 9]) with random-init weights.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl own|reference] [--config motion|scaled]

``--config scaled`` is BASELINE.json's config 4: the same classifier scaled to LSTM 9->1024 x 2 layers, T=512,
bf16 operands (fp32 master weights / accumulation / state), 64 samples per GPU (weak scaling: the persistent
recurrent kernels hold up to 128 rows), data-parallel with per-layer gradient buckets overlapped with the
backward pass (ops/layered_engine.py); its reference arm is the reference's own model and trainer under
``torch.autocast(bfloat16)`` + NCCL DistributedDataParallel.

N > 1 is launched by the driver under torchrun (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* env).
Rank 0 prints ONE JSON line.  Two measurements per run:

* ``value``   — device time: every step bracketed by CUDA events on the launching stream, a
  256 MiB L2 flush between steps (outside the brackets), summed over K steps, max over ranks.
  The step is fused forward -> fused head/CE/backward -> fused grad-reduce/all-reduce/Adam,
  3 launches of this repo's kernels; the batch is read from the device-resident dataset through
  the sampler's index slice by the kernels' own loaders.
* ``e2e``     — wall clock through the public API a user calls
  (``FusedTrainEngine.train_steps_dataset``): the dataset sits in pinned host memory; every step
  the sampler's index slice goes to the GPU, the gather kernel pulls that step's rows out of the
  pinned dataset over PCIe into a device batch (what the reference's DataLoader does with
  __getitem__ + collate + a host->device copy), and the step's loss is read back to the host;
  transfers run on a side stream under the previous step's compute.
* ``epoch``   — the BASELINE.md metric itself: wall clock of ONE epoch (6 912 samples: 4 full
  batches + the partial one) through ``Trainer.train(1)`` (src/motion/trainer/base.py:93-96), cold
  (first call, includes buffer allocation) and warm (second call); max over ranks.

``--impl reference`` runs the UNMODIFIED reference (installed at baseline/_ref) through its own
Trainer loop on the same GPUs (stock torch cuDNN LSTM + NCCL DistributedDataParallel); see
baseline/ref_adapter.py.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
# stdout carries exactly ONE JSON line.  Libraries write banners to file descriptor 1 behind Python's back (NCCL prints
# "NCCL version ..." there), so fd 1 is pointed at stderr for the whole run and the JSON line goes to a duplicate of
# the original stdout.
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
_REAL_STDOUT = os.fdopen(os.dup(1), "w")
os.dup2(2, 1)


def emit(obj) -> None:
    _REAL_STDOUT.write(json.dumps(obj) + "\n")
    _REAL_STDOUT.flush()

GLOBAL_BATCH, SEQ_LEN, FEATURES, HIDDEN, LAYERS, CLASSES, TRAIN_SIZE = 1440, 128, 9, 32, 2, 6, 6912
LR, SEED = 0.0025, 123456789
SCALED = dict(hidden=1024, layers=2, seq_len=512, per_gpu_batch=64, batches_per_epoch=8)
# best published number of the reference at batch 1440: 6912 samples / 3.595 s on its unnamed fast
# host (results_baseline_1.json results[0])
BASELINE_SAMPLES_PER_S = 6912 / 3.595
METRIC = "motion_lstm_ddp_train_samples_per_sec"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons, sampled while the benchmark runs (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.gpu = gpu_index
        self.rows = []  # (t, sm, max, {reason: bool})
        self.proc = None
        self.thread = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.gpu)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
        except OSError:
            return
        self.thread = threading.Thread(target=self._read, daemon=True)
        self.thread.start()

    def _read(self):
        for line in self.proc.stdout:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 9:
                continue
            try:
                sm, mx = float(f[1]), float(f[2])
            except ValueError:
                continue
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            self.rows.append((time.time(), sm, mx, {n: f[5 + i].lower().startswith("active") for i, n in enumerate(names)}))

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except subprocess.TimeoutExpired:
                self.proc.kill()

    def summary(self, windows):
        def inside(t):
            return any(a - 0.05 <= t <= b + 0.05 for a, b in windows)
        rows = [r for r in self.rows if inside(r[0])]
        scope = "timed_regions"
        if not rows:
            rows, scope = self.rows, "whole_benchmark_process(timed region shorter than the 100 ms sampling period)"
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0, "window": "nvidia-smi unavailable"}
        reasons = sorted({n for r in rows for n, on in r[3].items() if on})
        return {"sm_mhz": statistics.median(r[1] for r in rows), "sm_max_mhz": max(r[2] for r in rows),
                "reasons": reasons, "samples": len(rows), "window": scope}


def dist_setup(n_gpus: int):
    import torch
    if n_gpus > 1:
        from pytorch_distributed_rnn_b200.parallel import process_group as pg
        info = pg.init_distributed()
        assert info.world_size == n_gpus, f"--gpus {n_gpus} but WORLD_SIZE={info.world_size} (launch under torchrun)"
        return info.rank, info.world_size, info.device
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    return 0, 1, dev


def max_over_ranks(x: float, world: int, dev) -> float:
    import torch
    if world == 1:
        return x
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier_sync(world: int, dev):
    import torch
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize(dev)


def run_own(args):
    import torch
    from pytorch_distributed_rnn_b200 import MotionModel, ParamArena, _native
    from pytorch_distributed_rnn_b200.data import synthetic
    from pytorch_distributed_rnn_b200.data.sampler import epoch_indices
    from pytorch_distributed_rnn_b200.ops.engine import FusedTrainEngine

    rank, world, dev = dist_setup(args.gpus)
    lib = _native.lib()
    torch.manual_seed(SEED)
    model = MotionModel(FEATURES, HIDDEN, LAYERS, CLASSES).to(dev)
    comm, arena, comm_kind = None, None, "none"
    if world > 1:
        from pytorch_distributed_rnn_b200.parallel.comm import make_comm
        comm = make_comm()
        comm_kind = comm.kind
        arena = comm.make_arena(model) if hasattr(comm, "make_arena") else ParamArena(model)
        comm.broadcast_flat(arena.flat[:arena.numel], root=0)
    engine = FusedTrainEngine(model, lr=LR, arena=arena, comm=comm)
    per_rank = GLOBAL_BATCH // world
    K, W = args.steps, args.warmup

    # synthetic dataset of the reference's shape, device-resident (value) and pinned on the host (e2e)
    x_all, y_all = synthetic.make_split(TRAIN_SIZE, seed=0)
    feats, labels = x_all.to(dev), y_all.to(dev)
    full = TRAIN_SIZE // GLOBAL_BATCH  # 4 full global batches per epoch (the 5th, partial one is skipped)

    def step_indices(s):
        e, j = divmod(s, full)
        return epoch_indices(TRAIN_SIZE, world, rank, e)[j * per_rank:(j + 1) * per_rank]

    idx_dev = [step_indices(s).to(dev) for s in range(W + K)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    sp = _native.stream_ptr(dev)

    sampler = ClockSampler(dev.index or 0) if rank == 0 else None
    if sampler:
        sampler.start()

    # ---------------- device-timed value ----------------
    for s in range(W):
        engine.train_step_gather(feats, labels, idx_dev[s])
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    barrier_sync(world, dev)
    launches0 = _native.launches()
    t_a = time.time()
    for s in range(K):
        lib.b200rnn_l2_flush(flush.data_ptr(), flush.numel(), sp)
        ev[s][0].record()
        engine.train_step_gather(feats, labels, idx_dev[W + s])
        ev[s][1].record()
    barrier_sync(world, dev)
    t_b = time.time()
    launches = _native.launches() - launches0
    dev_ms = sum(a.elapsed_time(b) for a, b in ev)
    dev_ms = max_over_ranks(dev_ms, world, dev)
    last = engine.metrics[(engine.n_steps - 1) % engine.METRIC_RING].cpu()

    # ---------------- end-to-end through the public API ----------------
    # Every step's batch is ASSEMBLED on the host inside the timed region (rows of the dataset selected by the
    # sampler's indices into pinned staging memory), then copied to the device; its loss is read back.
    x_pin, y_pin = x_all.pin_memory(), y_all.pin_memory()       # the dataset, pinned once (as a DataLoader would pin batches)
    idx_host = [step_indices(s) for s in range(max(W, 3) + K)]  # the sampler's output: one index slice per step
    for _ in engine.train_steps_dataset(x_pin, y_pin, idx_host[:max(W, 3)]):
        pass
    barrier_sync(world, dev)
    t_c = time.time()
    t0 = time.perf_counter()
    n_results = 0
    for loss, correct in engine.train_steps_dataset(x_pin, y_pin, idx_host[max(W, 3):]):
        n_results += 1
    assert n_results == K
    barrier_sync(world, dev)
    e2e_s = time.perf_counter() - t0
    t_d = time.time()
    e2e_s = max_over_ranks(e2e_s, world, dev)
    h2d = world * per_rank * (SEQ_LEN * FEATURES * 4 + 8 + 8)   # rows + labels + the index slice
    d2h = world * 8

    # ---------------- one epoch through the trainer (the BASELINE.md metric) ----------------
    epoch = run_own_epoch(world, rank, dev, comm, x_all, y_all)

    clocks = None
    if sampler:
        sampler.stop()
        clocks = sampler.summary([(t_a, t_b), (t_c, t_d)])

    # a number produced while an in-kernel watchdog fired (a CTA gave up waiting for its producer) is garbage:
    # fail loudly instead of printing it (outside every timed region)
    _native.assert_device_healthy()

    if rank == 0:
        value = GLOBAL_BATCH * K / (dev_ms / 1e3)
        out = {
            "metric": METRIC, "value": value, "unit": "samples/s", "n_gpus": world, "steps": K, "warmup": W,
            "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": value / BASELINE_SAMPLES_PER_S, "dtype": "fp32", "data": "synthetic",
            "impl": "own",
            "config": {"model": "motion LSTM 9->32 x2 layers, T=128, Linear 32->6, CE, Adam lr 0.0025",
                       "global_batch": GLOBAL_BATCH, "seq_len": SEQ_LEN, "parallelism": f"dp{world}",
                       "per_rank_batch": per_rank, "comm": comm_kind,
                       "l2": "256 MiB L2 flush between timed steps, outside the CUDA-event brackets",
                       "timing": "per-step CUDA events on the launching stream, summed, max over ranks",
                       "baseline_number": "1922.7 samples/s = 6912/3.595 s, reference local bs1440 on its fast host"},
            "e2e": {"value": GLOBAL_BATCH * K / e2e_s, "unit": "samples/s", "ms_per_step": 1e3 * e2e_s / K,
                    "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "api": "FusedTrainEngine.train_steps_dataset(pinned_features, pinned_labels, index_slices) -> yields (loss, correct) per step",
                    "inputs": "every step, inside the timed region: the sampler's index slice goes host -> device, the gather kernel "
                              "pulls that batch's rows from the PINNED HOST dataset into device memory (side stream, overlapping the "
                              "previous step), and the step's loss is read device -> host (one-step lag); nothing is pre-collated"},
            "epoch": epoch,
            "gpu_launches": launches,
            "launches_per_step": launches / K,
            "clocks": clocks,
            "final_loss": float(last[0]) / per_rank,
        }
        emit(out)
    if world > 1:
        import torch.distributed as dist
        barrier_sync(world, dev)
        dist.destroy_process_group()


def run_own_scaled(args):
    """BASELINE.json config 4 (see the module docstring).  Same three measurements as the motion config."""
    import torch
    from pytorch_distributed_rnn_b200 import MotionModel, ParamArena, _native
    from pytorch_distributed_rnn_b200.data.sampler import epoch_indices
    from pytorch_distributed_rnn_b200.ops.layered_engine import LayeredTrainEngine

    rank, world, dev = dist_setup(args.gpus)
    lib = _native.lib()
    H, L, T, per_rank = SCALED["hidden"], SCALED["layers"], SCALED["seq_len"], SCALED["per_gpu_batch"]
    gb = per_rank * world
    n_train = gb * SCALED["batches_per_epoch"]
    torch.manual_seed(SEED)
    model = MotionModel(FEATURES, H, L, CLASSES).to(dev)
    comm, arena, comm_kind = None, None, "none"
    if world > 1:
        from pytorch_distributed_rnn_b200.parallel.comm import make_comm
        comm = make_comm()
        comm_kind = comm.kind
        arena = comm.make_arena(model) if hasattr(comm, "make_arena") else ParamArena(model)
        comm.broadcast_flat(arena.flat[:arena.numel], root=0)
    engine = LayeredTrainEngine(model, lr=LR, arena=arena, comm=comm)
    K, W = args.steps, args.warmup
    g = torch.Generator().manual_seed(0)
    y_all = torch.randint(0, CLASSES, (n_train, 1), generator=g)
    x_all = torch.randn(n_train, T, FEATURES, generator=g)
    feats, labels = x_all.to(dev), y_all.to(dev)
    full = SCALED["batches_per_epoch"]

    def step_indices(s):
        e, j = divmod(s, full)
        return epoch_indices(n_train, world, rank, e)[j * per_rank:(j + 1) * per_rank]

    idx_dev = [step_indices(s).to(dev) for s in range(W + K)]
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    sp = _native.stream_ptr(dev)
    sampler = ClockSampler(dev.index or 0) if rank == 0 else None
    if sampler:
        sampler.start()
    for s in range(W):
        engine.train_step_gather(feats, labels, idx_dev[s])
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
    barrier_sync(world, dev)
    launches0 = _native.launches()
    t_a = time.time()
    for s in range(K):
        lib.b200rnn_l2_flush(flush.data_ptr(), flush.numel(), sp)
        ev[s][0].record()
        engine.train_step_gather(feats, labels, idx_dev[W + s])
        ev[s][1].record()
    barrier_sync(world, dev)
    t_b = time.time()
    launches = _native.launches() - launches0
    dev_ms = max_over_ranks(sum(a.elapsed_time(b) for a, b in ev), world, dev)
    last = engine.metrics[(engine.n_steps - 1) % engine.METRIC_RING].cpu()

    x_pin, y_pin = x_all.pin_memory(), y_all.pin_memory()
    idx_host = [step_indices(s) for s in range(max(W, 3) + K)]
    for _ in engine.train_steps_dataset(x_pin, y_pin, idx_host[:max(W, 3)]):
        pass
    barrier_sync(world, dev)
    t_c = time.time()
    t0 = time.perf_counter()
    n_results = sum(1 for _ in engine.train_steps_dataset(x_pin, y_pin, idx_host[max(W, 3):]))
    assert n_results == K
    barrier_sync(world, dev)
    e2e_s = max_over_ranks(time.perf_counter() - t0, world, dev)
    t_d = time.time()
    clocks = None
    if sampler:
        sampler.stop()
        clocks = sampler.summary([(t_a, t_b), (t_c, t_d)])
    _native.assert_device_healthy()
    if comm is not None and hasattr(comm, "check"):
        comm.check()
    assert engine.reducer_status() == 0, "a reducer trigger wait timed out"
    # one more step, outside every timed region, with %globaltimer stamps: when each gradient bucket's trigger fired,
    # when its collective / optimizer finished, and when each layer's backward began and ended
    engine.timeline = True
    engine.train_step_gather(feats, labels, idx_dev[-1])
    timeline = engine.last_timeline()
    engine.timeline = False
    if rank == 0:
        value = gb * K / (dev_ms / 1e3)
        out = {
            "metric": "scaled_lstm_ddp_train_samples_per_sec", "value": value, "unit": "samples/s", "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": dev_ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": "own",
            "config": {"model": f"LSTM 9->{H} x{L} layers, T={T}, Linear {H}->6, CE, Adam lr 0.0025 (BASELINE.json config 4)",
                       "global_batch": gb, "seq_len": T, "parallelism": f"dp{world}", "per_rank_batch": per_rank,
                       "comm": comm_kind, "precision": "bf16 GEMM operands, fp32 master weights / accumulation / cell state",
                       "buckets": [{"name": n, "bytes": 4 * (hi - lo)} for n, lo, hi in engine.buckets],
                       "l2": "256 MiB L2 flush between timed steps, outside the CUDA-event brackets",
                       "timing": "per-step CUDA events on the launching stream, summed, max over ranks"},
            "e2e": {"value": gb * K / e2e_s, "unit": "samples/s", "ms_per_step": 1e3 * e2e_s / K,
                    "h2d_bytes_per_step": world * per_rank * (T * FEATURES * 4 + 8 + 8), "d2h_bytes_per_step": world * 8,
                    "api": "LayeredTrainEngine.train_steps_dataset(pinned_features, pinned_labels, index_slices) -> yields (loss, correct) per step",
                    "inputs": "every step, inside the timed region: index slice host -> device, gather of the batch's rows from the pinned "
                              "host dataset by the GPU, D2H loss read"},
            "gpu_launches": launches, "launches_per_step": launches / K, "clocks": clocks,
            "overlap_timeline_us": timeline,
            "final_loss": float(last[0]) / per_rank,
        }
        emit(out)
    if world > 1:
        import torch.distributed as dist
        barrier_sync(world, dev)
        dist.destroy_process_group()


def run_own_epoch(world, rank, dev, comm, x_all, y_all):
    """Wall clock of one epoch (5 steps incl. the partial 1 152-sample batch) through Trainer.train(1): the public
    training entry point, dataset device-resident (uploaded by the trainer's constructor, outside the clock — the
    reference loads its dataset before the clock starts too, trainer/__init__.py:24-27 vs base.py:93)."""
    import logging
    import torch
    from pytorch_distributed_rnn_b200 import MotionModel
    from pytorch_distributed_rnn_b200.data.dataset import MotionDataset
    from pytorch_distributed_rnn_b200.trainer.base import Trainer
    logging.getLogger().setLevel("WARNING")
    torch.manual_seed(SEED)
    model = MotionModel(FEATURES, HIDDEN, LAYERS, CLASSES)
    train = MotionDataset(x_all, y_all)
    if world > 1:
        from pytorch_distributed_rnn_b200.trainer.ddp import DDPTrainer
        trainer = DDPTrainer(model, train, GLOBAL_BATCH, LR, device=dev, comm=comm)
    else:
        trainer = Trainer(model, train, GLOBAL_BATCH, LR, device=dev)
    times = []
    for _ in range(2):
        barrier_sync(world, dev)
        t0 = time.perf_counter()
        trainer.train(1)
        torch.cuda.synchronize(dev)
        times.append(max_over_ranks(time.perf_counter() - t0, world, dev))
    return {"first_epoch_s": times[0], "epoch_s": times[1], "samples": TRAIN_SIZE, "steps": 5,
            "samples_per_s": TRAIN_SIZE / times[1],
            "api": "Trainer.train(1) / DDPTrainer.train(1), wall clock, max over ranks"}


def run_reference(args):
    try:
        from baseline import ref_adapter
    except Exception as e:  # noqa: BLE001
        emit({"impl": "reference", "unavailable": f"adapter import failed: {type(e).__name__}: {e}"})
        return
    why = ref_adapter.unavailable_reason()
    if why:
        rank = int(os.environ.get("RANK", "0"))
        if rank == 0:
            emit({"impl": "reference", "unavailable": why})
        return
    rank, world, dev = dist_setup_reference(args.gpus)
    sampler = ClockSampler(dev.index or 0) if rank == 0 else None
    if sampler:
        sampler.start()
    scaled = args.config == "scaled"
    gb = SCALED["per_gpu_batch"] * world if scaled else GLOBAL_BATCH
    cfg = dict(global_batch=gb, seq_len=SEQ_LEN, features=FEATURES, hidden=HIDDEN, layers=LAYERS, classes=CLASSES,
               lr=LR, seed=SEED, train_size=TRAIN_SIZE, autocast_bf16=False, epoch=True)
    if scaled:
        cfg.update(seq_len=SCALED["seq_len"], hidden=SCALED["hidden"], layers=SCALED["layers"],
                   train_size=gb * SCALED["batches_per_epoch"], autocast_bf16=True, epoch=False)
    res = ref_adapter.run(args.gpus, args.steps, args.warmup, rank, world, dev, cfg)
    clocks = None
    if sampler:
        sampler.stop()
        clocks = sampler.summary([res["window"]])
    dev_s = max_over_ranks(res["device_s"], world, dev)
    wall_s = max_over_ranks(res["wall_s"], world, dev)
    if rank == 0:
        K = args.steps
        value = gb * K / dev_s
        out = {
            "metric": "scaled_lstm_ddp_train_samples_per_sec" if scaled else METRIC, "value": value, "unit": "samples/s",
            "n_gpus": world, "steps": K, "warmup": args.warmup,
            "ms_per_step": 1e3 * dev_s / K, "higher_is_better": True, "scaling": "weak" if scaled else "strong",
            "vs_baseline": None if scaled else value / BASELINE_SAMPLES_PER_S, "dtype": "bf16" if scaled else "fp32",
            "data": "synthetic", "impl": "reference",
            "config": {"model": f"reference MotionModel (nn.LSTM 9->{cfg['hidden']} x{cfg['layers']}, T={cfg['seq_len']}, "
                                f"Linear {cfg['hidden']}->6), CE, Adam lr 0.0025" + (", torch.autocast(bfloat16)" if scaled else ""),
                       "global_batch": gb, "seq_len": cfg["seq_len"], "parallelism": f"dp{world}",
                       "per_rank_batch": gb // world,
                       "path": res["path"],
                       "timing": "CUDA events around the reference's own _train_step loop (K steps), max over ranks"},
            "e2e": {"value": gb * K / wall_s, "unit": "samples/s", "ms_per_step": 1e3 * wall_s / K,
                    "h2d_bytes_per_step": res["h2d_bytes_per_step"], "d2h_bytes_per_step": res["d2h_bytes_per_step"]},
            "epoch": res.get("epoch"),
            "gpu_launches": None,
            "clocks": clocks,
            "final_loss": res.get("final_loss"),
        }
        emit(out)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()


def dist_setup_reference(n_gpus: int):
    """Same rendezvous as the own arm, but with torch only (none of this repo's engine on the reference path)."""
    import torch
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    if n_gpus > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=dev)
        assert dist.get_world_size() == n_gpus
        return dist.get_rank(), dist.get_world_size(), dev
    return 0, 1, dev


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="own", choices=["own", "reference"])
    ap.add_argument("--config", default="motion", choices=["motion", "scaled"])
    args = ap.parse_args()
    if args.config == "scaled" and "--steps" not in sys.argv:
        args.steps, args.warmup = 10, 3   # ~25 ms per step
    if args.warmup < 3:
        args.warmup = 3  # timing rule: at least 3 warm-up steps
    import torch
    if not torch.cuda.is_available():
        emit({"impl": args.impl, "unavailable": "no CUDA device visible"})
        return
    if args.impl == "reference":
        run_reference(args)
    elif args.config == "scaled":
        run_own_scaled(args)
    else:
        run_own(args)


if __name__ == "__main__":
    main()
