"""Reference arm of bench.py — NOT product code.

Runs the UNMODIFIED reference (jkhlr/pytorch-distributed-rnn, installed byte-for-byte under
``baseline/_ref`` by ``baseline/install_reference.sh``) through its own classes and its own hot
loop: ``MotionModel`` (src/motion/model.py), ``MotionDataset`` (dataset.py), ``Trainer`` /
``DistributedTrainer`` (trainer/base.py, trainer/distributed.py) including the stock
``DataLoader`` + ``DistributedSampler`` + ``Adam`` they construct, and ``Trainer._train_step``.
None of this repo's models, kernels or engine is imported here.

The reference has no GPU path (torch 1.4 CPU build, 'mpi' backend hard-coded at
trainer/ddp.py:18, tensors never moved to a device), so the harness supplies only what a user
porting it to a GPU box must: (a) stand-ins for the two modules it imports unconditionally but
that are not installable offline (``horovod.torch`` — unused by the DDP path — and
``memory_profiler.memory_usage``), (b) a DistributedTrainer subclass whose constructor mirrors
DDPTrainer's but says ``init_process_group('nccl')`` / ``model.cuda()`` /
``DistributedDataParallel(model, device_ids=[gpu])`` (stock torch => cuDNN LSTM + NCCL
all-reduce), and (c) a loader wrapper that moves each collated batch to the GPU and limits the
number of batches so exactly K steps are timed.
"""
from __future__ import annotations

import logging
import sys
import time
import types
from pathlib import Path

REF_ROOT = Path(__file__).resolve().parent / "_ref"
REF_MOTION = REF_ROOT / "motion"


def unavailable_reason():
    """None when the reference is importable; installs it first if baseline/_ref is missing."""
    from baseline import ensure_reference
    try:
        ensure_reference()
    except Exception as e:  # noqa: BLE001
        return f"reference not installable under baseline/_ref: {type(e).__name__}: {str(e)[:300]}"
    return None


def _install_stubs():
    if "memory_profiler" not in sys.modules:
        mp = types.ModuleType("memory_profiler")

        def memory_usage(proc, *a, **kw):
            import resource
            f, args, kwargs = proc
            f(*args, **kwargs)
            return [resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1024.0]
        mp.memory_usage = memory_usage
        sys.modules["memory_profiler"] = mp
    if "horovod" not in sys.modules:
        hv = types.ModuleType("horovod")
        hvt = types.ModuleType("horovod.torch")
        hv.torch = hvt
        sys.modules["horovod"] = hv
        sys.modules["horovod.torch"] = hvt


class _DeviceLoader:
    """Yields exactly ``n_batches`` batches from the reference's own DataLoader (re-seeding its
    sampler per epoch like Trainer.train does), moved to the GPU."""

    def __init__(self, loader, sampler, device, n_batches, counter):
        self.loader, self.sampler, self.device, self.n = loader, sampler, device, n_batches
        self.dataset = loader.dataset
        self.counter = counter

    def __len__(self):
        return self.n

    def __iter__(self):
        done, epoch = 0, 0
        while done < self.n:
            self.sampler.set_epoch(epoch)
            for data, labels in self.loader:
                if data.shape[0] != self.loader.batch_size:
                    continue  # skip the partial tail batch so every timed step is a full batch
                self.counter["h2d"] = data.numel() * data.element_size() + labels.numel() * labels.element_size()
                yield data.to(self.device), labels.to(self.device)
                done += 1
                if done >= self.n:
                    return
            epoch += 1


class _EpochLoader:
    """The reference's own DataLoader, every batch (incl. the partial last one) moved to the GPU."""

    def __init__(self, loader, device):
        self.loader, self.device, self.dataset = loader, device, loader.dataset

    def __len__(self):
        return len(self.loader)

    def __iter__(self):
        for data, labels in self.loader:
            yield data.to(self.device), labels.to(self.device)


def run(n_gpus, steps, warmup, rank, world, device, cfg):
    import torch
    _install_stubs()
    if str(REF_MOTION) not in sys.path:
        sys.path.insert(0, str(REF_MOTION))
    from dataset import MotionDataset            # reference
    from model import MotionModel                # reference
    from trainer.base import Trainer             # reference
    from trainer.distributed import DistributedTrainer  # reference
    from trainer.formatter import TrainingMessageFormatter  # reference

    logging.getLogger().setLevel("WARNING")  # the reference's --log flag; same level as the own arm
    torch.manual_seed(cfg["seed"])
    # synthetic data of the reference's shape (same generator as the own arm, plain torch)
    g = torch.Generator().manual_seed(0)
    n = cfg.get("train_size", 6912)
    y = torch.randint(0, cfg["classes"], (n, 1), generator=g)
    x = torch.randn(n, cfg["seq_len"], cfg["features"], generator=g)
    train = MotionDataset(x, y)
    model = MotionModel(cfg["features"], cfg["hidden"], cfg["layers"], cfg["classes"])

    if world > 1:
        from torch.nn.parallel.distributed import DistributedDataParallel

        class NcclDDPTrainer(DistributedTrainer):  # DDPTrainer (trainer/ddp.py) with 'mpi' -> 'nccl' + .cuda()
            def __init__(self, model, training_set, batch_size, learning_rate):
                m = DistributedDataParallel(model.to(device), device_ids=[device.index])
                super().__init__(rank=rank, world_size=world, model=m, training_set=training_set,
                                 batch_size=batch_size, learning_rate=learning_rate)
        trainer = NcclDDPTrainer(model, train, cfg["global_batch"], cfg["lr"])
        path = "reference DistributedTrainer + torch DistributedDataParallel(NCCL) + cuDNN nn.LSTM"
    else:
        trainer = Trainer(model.to(device), train, cfg["global_batch"], cfg["lr"])
        path = "reference Trainer (local) + cuDNN nn.LSTM"
    formatter = TrainingMessageFormatter(1, rank)
    counter = {"h2d": 0}
    base_loader, sampler = trainer.train_loader, trainer.sampler

    import contextlib
    # scaled config: the reference's loop under bf16 autocast (what a user of the reference would switch on for
    # a bf16 run; the loop itself is untouched)
    amp = (lambda: torch.autocast("cuda", dtype=torch.bfloat16)) if cfg.get("autocast_bf16") else contextlib.nullcontext
    if cfg.get("autocast_bf16"):
        path += " under torch.autocast(bfloat16)"

    def run_steps(k):
        trainer.train_loader = _DeviceLoader(base_loader, sampler, device, k, counter)
        with amp():
            return trainer._train_step(formatter)  # the reference's own hot loop (trainer/base.py:103-130)

    run_steps(max(warmup, 3))
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize(device)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t_a = time.time()
    t0 = time.perf_counter()
    e0.record()
    loss, _ = run_steps(steps)
    e1.record()
    torch.cuda.synchronize(device)
    if world > 1:
        torch.distributed.barrier()
    wall = time.perf_counter() - t0
    t_b = time.time()
    # one epoch through the reference's own Trainer.train(1) (trainer/base.py:65-101): the BASELINE.md metric
    trainer.train_loader = _EpochLoader(base_loader, device)
    ep = []
    for _ in range(2 if cfg.get("epoch", True) else 0):
        if world > 1:
            torch.distributed.barrier()
        torch.cuda.synchronize(device)
        te = time.perf_counter()
        trainer.train(1)
        torch.cuda.synchronize(device)
        ep.append(time.perf_counter() - te)
    if world > 1 and ep:
        t = torch.tensor(ep, dtype=torch.float64, device=device)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ep = [float(v) for v in t.tolist()]
    epoch = None
    if ep:
        epoch = {"first_epoch_s": ep[0], "epoch_s": ep[1], "samples": n, "steps": len(base_loader),
                 "samples_per_s": n / ep[1], "api": "reference Trainer.train(1), wall clock, max over ranks"}
    return {"epoch": epoch,"device_s": e0.elapsed_time(e1) / 1e3, "wall_s": wall, "window": (t_a, t_b), "path": path,
            "h2d_bytes_per_step": counter["h2d"] * world, "d2h_bytes_per_step": 8 * world,
            "final_loss": None}
