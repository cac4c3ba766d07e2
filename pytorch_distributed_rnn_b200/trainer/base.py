"""Local trainer — the training loop all trainers share.

Behavioural parity with src/motion/trainer/base.py: epoch-seeded sharded sampling even for
local runs (:22-26,74-75), train batches of ``batch_size`` and ONE full-dataset batch for
validation / test (:32-39,53-54), per-batch progress lines, per-epoch validation with a
``best-model.pt`` checkpoint on improvement (:80-91,164-177), the whole loop timed and its
peak memory reported in the performance line (:93-96), a final test evaluation (:98-99) and
``train()`` returning ``(model.eval(), train_history, validation_history)`` with
``train_history[e] = sum(batch mean losses) / len(dataset)`` (:128, quirk Q5 kept).

What is different is underneath.  With the model on a CUDA device the loop drives
:class:`~..ops.engine.FusedTrainEngine`: the dataset lives on the GPU, a batch is a K11
gather over the epoch's index vector, and forward + loss + backward + optimizer are three
kernel launches with no host synchronisation; progress lines are rendered from a device
ring buffer (per step, or once per epoch with ``deferred_logging=True``).  On CPU the same
loop runs plain torch autograd — the plumbing/oracle configuration.
"""
from __future__ import annotations

import logging
import os
import resource
import time
from pathlib import Path
from typing import List, Optional

import torch
from torch import nn

from ..data.sampler import ShardedEpochSampler
from ..ops.optim import FusedAdam
from ..utils.arena import ParamArena
from .formatter import TrainingMessageFormatter


class _Range:
    """NVTX range when B200RNN_NVTX=1 (for nsys / ncu timelines); free otherwise.  The reference has
    no tracing beyond a wall clock and memory_profiler."""
    on = os.environ.get("B200RNN_NVTX", "0") == "1"

    def __init__(self, name: str):
        self.name = name

    def __enter__(self):
        if self.on and torch.cuda.is_available():
            torch.cuda.nvtx.range_push(self.name)

    def __exit__(self, *a):
        if self.on and torch.cuda.is_available():
            torch.cuda.nvtx.range_pop()


def peak_rss_mib() -> float:
    """Peak resident set size of this process in MiB (what memory_profiler reports, base.py:94)."""
    return resource.getrusage(resource.RUSAGE_SELF).ru_maxrss / 1024.0


class Trainer:
    loss_fn = nn.CrossEntropyLoss()

    def __init__(self, model: nn.Module, training_set, batch_size: int, learning_rate: float,
                 validation_set=None, test_set=None, checkpoint_dir=None, sampler=None,
                 device: Optional[torch.device] = None, deferred_logging: bool = False,
                 save_every: int = 0, resume=None):
        self.rank = getattr(self, "rank", 0)
        self.world_size = getattr(self, "world_size", 1)
        self.device = torch.device(device) if device is not None else self._default_device()
        self.model = self._place_model(model)
        self.checkpoint_dir = Path(checkpoint_dir) if checkpoint_dir is not None else None
        self.deferred_logging = deferred_logging
        self.save_every = save_every
        self.training_set = self._place_dataset(training_set)
        self.validation_set = self._place_dataset(validation_set)
        self.test_set = self._place_dataset(test_set)
        self.sampler = sampler or ShardedEpochSampler(len(training_set), num_replicas=1, rank=0)
        self.batch_size = self._per_rank_batch(batch_size)
        self.arena = self._get_arena()
        self.optimizer = self._get_optimizer(self.model, learning_rate)
        self.engine = self._get_engine()
        self.start_epoch = 0
        self.last_duration = None
        self.last_device_seconds = None
        if resume is not None:
            self.load_checkpoint(resume)

    # ---- construction hooks (overridden by the distributed trainers) ----
    def _default_device(self) -> torch.device:
        return torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu")

    def _place_model(self, model: nn.Module) -> nn.Module:
        return model.to(self.device)

    def _place_dataset(self, ds):
        if ds is None:
            return None
        return ds.to(self.device) if hasattr(ds, "to") else ds

    def _per_rank_batch(self, batch_size: int) -> int:
        return batch_size

    def _core_model(self) -> nn.Module:
        return self.model

    def _get_arena(self) -> ParamArena:
        return ParamArena(self._core_model())

    def _get_optimizer(self, model, lr):
        return FusedAdam(self.arena, lr=lr)

    def _get_comm(self):
        return None

    def _get_engine(self):
        if self.device.type != "cuda":
            return None
        from ..models.motion import MotionModel
        from ..ops.engine import FusedTrainEngine
        core = self._core_model()
        if not isinstance(core, MotionModel):
            return None
        try:
            return FusedTrainEngine(core, optimizer=self._flat_optimizer(), arena=self.arena, comm=self._get_comm())
        except NotImplementedError:
            pass
        try:   # large-hidden family: tcgen05 GEMMs + persistent recurrence, per-layer gradient buckets
            from ..ops.layered_engine import LayeredTrainEngine
            return LayeredTrainEngine(core, optimizer=self._flat_optimizer(), arena=self.arena, comm=self._get_comm())
        except NotImplementedError as e:
            logging.warning("no fused engine for this shape (%s); using the composite torch path", e)
            return None

    def _flat_optimizer(self):
        return getattr(self.optimizer, "_opt", self.optimizer)

    def _get_formatter(self, epochs):
        return TrainingMessageFormatter(epochs)

    # ---- the loop ----
    def train(self, epochs: int):
        training_history: List[float] = []
        validation_history: List[float] = []
        formatter = self._get_formatter(epochs)
        cuda = self.device.type == "cuda"
        if cuda:
            torch.cuda.reset_peak_memory_stats(self.device)
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize(self.device)
            ev0.record()
        start = time.perf_counter()
        best_loss = None
        for epoch in range(self.start_epoch, epochs):
            if self.sampler is not None:
                self.sampler.set_epoch(epoch)
            logging.info(formatter.epoch_start_message(epoch))
            with _Range(f"epoch{epoch}/train"):
                train_loss, _ = self._train_step(formatter)
            training_history.append(train_loss)
            if self.validation_set is not None:
                with _Range(f"epoch{epoch}/validate"):
                    validation_loss, _ = self._evaluate(self.validation_set, formatter, epoch)
                validation_history.append(validation_loss)
                if best_loss is None or best_loss > validation_loss:
                    logging.info(f"New best model in epoch {epoch + 1}")
                    best_loss = validation_loss
                    self._save_checkpoint(epoch, validation_loss, best=True)
            if self.save_every and (epoch + 1) % self.save_every == 0:
                self._save_checkpoint(epoch, training_history[-1], best=False)
            # a fired watchdog (stuck peer, lost CTA) means this epoch trained on garbage: stop now, not after
            # the last epoch; then keep the ranks together after rank-0-only work (validation, checkpoints)
            self._health_check()
            self._epoch_barrier()
        if cuda:
            ev1.record()
            torch.cuda.synchronize(self.device)
            self.last_device_seconds = ev0.elapsed_time(ev1) / 1e3
        duration = time.perf_counter() - start
        self.last_duration = duration
        self._health_check()
        logging.info(formatter.performance_message(peak_rss_mib(), duration))
        if cuda:
            logging.info("Rank: {:02d}   Device time: {:.6f} s, peak device memory: {:.1f} MiB".format(
                self.rank, self.last_device_seconds, torch.cuda.max_memory_allocated(self.device) / 2 ** 20))
        if self.test_set is not None:
            self._evaluate(self.test_set, formatter)
        return self.model.eval(), training_history, validation_history

    def _health_check(self) -> None:
        """Raise if an in-kernel watchdog fired: the comm kernels' flag waits (dead / stuck peer) or the compute
        kernels' own bounded waits (TMA / MMA / layer wavefront).  One small D2H read each."""
        comm = self._get_comm()
        if comm is not None and hasattr(comm, "check"):
            comm.check()
        if self.device.type == "cuda":
            from .. import _native
            _native.assert_device_healthy()
        status = getattr(self.engine, "reducer_status", None)
        if status is not None and status() != 0:
            raise RuntimeError(f"gradient reducer: a bucket's device-side trigger wait timed out (status {status():#x})")

    def _epoch_barrier(self) -> None:
        """Hook: distributed trainers wait here for rank 0's validation / checkpoint of this epoch."""

    def _epoch_batches(self):
        idx = self.sampler.indices() if hasattr(self.sampler, "indices") else torch.tensor(list(iter(self.sampler)))
        return [idx[i:i + self.batch_size] for i in range(0, idx.numel(), self.batch_size)]

    def _train_step(self, formatter):
        self.model.train()
        batches = self._epoch_batches()
        n_batches = len(batches)
        ds = self.training_set
        total_loss = 0.0
        total_correct = 0
        pending = []  # (batch_idx, n, metrics view) awaiting a host read

        def flush():
            nonlocal total_loss, total_correct
            if not pending:
                return
            vals = torch.stack([m for _, _, m in pending]).cpu()  # one D2H for all pending steps
            for (bi, n, _), v in zip(pending, vals):
                loss, correct = float(v[0]) / n, int(round(float(v[1])))
                total_loss += loss
                total_correct += correct
                logging.info(formatter.train_progress_message(
                    batch_idx=bi, batches=n_batches, training_examples=n, correct=correct, loss=loss))
            pending.clear()

        # the epoch's index vector goes to the device once (27 KB for the reference's dataset); batches are slices of it
        idx_dev, off = None, 0
        if self.engine is not None and batches:
            idx_dev = torch.cat(batches).to(self.device, non_blocking=True)
        # per-step progress lines need a host read per step (the reference reads loss.item() twice per step); when
        # they are not going to be emitted (log level above INFO) the metrics stay in the device ring until the epoch ends
        per_step = not self.deferred_logging and logging.getLogger().isEnabledFor(logging.INFO)
        for batch_idx, idx in enumerate(batches):
            n = idx.numel()
            if self.engine is not None:
                m = self.engine.train_step_gather(ds.features, ds.labels, idx_dev[off:off + n])
                off += n
                # a ring slot is only reused after METRIC_RING steps: no copy needed for shorter epochs
                keep = per_step or n_batches <= getattr(self.engine, "METRIC_RING", 0)
                pending.append((batch_idx, n, m if keep else m.clone()))
            else:
                pending.append((batch_idx, n, self._torch_train_step(ds, idx)))
            if per_step:
                flush()
        flush()
        return total_loss / len(ds), total_correct / len(ds)

    def _torch_train_step(self, ds, idx) -> torch.Tensor:
        """Generic autograd step (CPU plumbing path / shapes outside the fused families)."""
        idx = idx.to(ds.features.device)
        data = ds.features[idx].to(self.device)
        labels = ds.labels[idx].to(self.device).long().reshape(-1)
        self.optimizer.zero_grad()
        output = self.model(data)
        loss = self.loss_fn(output, labels)
        correct = (torch.argmax(output, dim=1) == labels).sum()
        loss.backward()
        self._sync_gradients()
        self.optimizer.step()
        return torch.stack([loss.detach() * labels.numel(), correct.to(loss.dtype)])

    def _sync_gradients(self) -> None:
        """Hook for trainers whose gradient averaging is not triggered by backward()/step() themselves."""

    def _evaluate(self, dataset, formatter, epoch=None):
        self.model.eval()
        n = len(dataset)
        with torch.no_grad():
            if self.engine is not None:
                v = self.engine.evaluate(dataset.features, dataset.labels).cpu()
                eval_loss, total_correct = float(v[0]) / n, int(round(float(v[1])))
            else:
                data = dataset.features.to(self.device)
                labels = dataset.labels.to(self.device).long().reshape(-1)
                output = self._core_model()(data)
                eval_loss = self.loss_fn(output, labels).item()
                total_correct = int((torch.argmax(output, dim=1) == labels).sum())
        accuracy = float(total_correct) / n
        logging.info(formatter.evaluation_message(accuracy, n, epoch, eval_loss, total_correct))
        return eval_loss, accuracy

    # ---- checkpoints ----
    def _save_checkpoint(self, epoch, loss, best=False):
        if self.checkpoint_dir is None:
            return
        self.checkpoint_dir.mkdir(parents=True, exist_ok=True)
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        checkpoint = {
            "epoch": epoch + 1,
            "model_state": {k: v.detach().cpu().clone() for k, v in self.model.state_dict().items()},
            "optimizer_state": self.optimizer.state_dict(),
            "loss": loss,
        }
        name = "best-model.pt" if best else f"checkpoint-epoch-{epoch + 1}.pt"
        torch.save(checkpoint, self.checkpoint_dir / name)

    def load_checkpoint(self, path) -> dict:
        """Extension (the reference has no resume path): accepts prefixed and un-prefixed keys."""
        ck = torch.load(path, map_location="cpu", weights_only=True)   # tensors + plain containers only
        state = ck["model_state"]
        own = self.model.state_dict()
        if any(k.startswith("module.") for k in state) and not any(k.startswith("module.") for k in own):
            state = {k[len("module."):]: v for k, v in state.items()}
        elif any(k.startswith("module.") for k in own) and not any(k.startswith("module.") for k in state):
            state = {"module." + k: v for k, v in state.items()}
        self.model.load_state_dict(state)
        self.arena.rebind()
        if ck.get("optimizer_state") is not None:
            self.optimizer.load_state_dict(ck["optimizer_state"])
        self.start_epoch = int(ck.get("epoch", 0))
        return ck
