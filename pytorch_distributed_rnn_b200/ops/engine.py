"""FusedTrainEngine — the whole training step of the motion classifier as 3 kernel launches.

(Exactly three in steady state, at every world size: the batch is read through the sampler's index
slice by the kernels' own loaders, the layer-wavefront sentinels are put back by their consumers,
and the reduction of the per-CTA gradient rows is the first phase of the update / all-reduce kernel.)

The reference's hot loop (src/motion/trainer/base.py:103-130) is, per batch:
``zero_grad -> model(data) -> CrossEntropyLoss -> argmax/eq/sum -> backward -> Adam.step``
— hundreds of ATen kernel launches, two host syncs (``loss.item()`` twice) and, under DDP,
a blocking all-reduce inside ``backward``.  Here the same step is:

1. ``b200rnn_lstm_fwd``   — all layers, all timesteps, one persistent launch (K2);
2. ``b200rnn_lstm_bwd``   — head + CE + correct-count + BPTT + dW accumulation (K5+K3+K4);
3. ``b200rnn_fused_update`` (1 GPU) or ``b200rnn_comm_allreduce_update`` (N GPUs) —
   deterministic reduction of the per-CTA gradient rows, the in-kernel NVLink all-reduce,
   gradient averaging and the Adam/SGD step (K6+K7);

The K11 gather of a device-resident dataset is folded into launches 1 and 2 (``train_step_gather``).
Nothing in the step synchronises with the host; loss / correct-count land in a device ring
buffer that the trainer reads when it wants to log.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, Optional, Tuple

import torch

from .. import _native
from .._native import ModelDesc
from ..models.motion import MotionModel
from ..utils.arena import ParamArena
from .optim import FusedAdam, _FlatOptimizer
from .rnn_function import CELL_ID


class HostStepsMixin:
    """End-to-end stepping from HOST batches for an engine that provides ``train_step(x_dev, y_dev)``, ``device``,
    ``_copy_stream`` and ``_host_metrics`` (shared by the small-hidden and the layered large-hidden engines)."""

    def train_steps_dataset(self, feats_pinned: torch.Tensor, labels_pinned: torch.Tensor, index_batches):
        """End-to-end training from a dataset that lives in PINNED HOST memory, given the sampler's index slices
        (an iterable of int64 host tensors, one per step) — the DataLoader + DistributedSampler of the reference
        (src/motion/trainer/base.py:56-60,107) without a host-side collate: per step the index slice goes to the
        device (a few KB) and the K11 gather kernel pulls that step's rows straight out of the pinned dataset over
        PCIe into a device batch, on a side stream while the previous step computes; the step's loss is read back
        with a one-step lag.  Every step still moves its whole input host -> device and its result device -> host.
        Generator: yields ``(mean loss, correct count)`` per step, in order."""
        return _train_steps_dataset(self, feats_pinned, labels_pinned, index_batches)

    def train_step_host(self, x_host: torch.Tensor, y_host: torch.Tensor) -> Tuple[float, int]:
        """End-to-end step from host memory: H2D copy of the batch (pass pinned tensors for an async
        DMA), the 3-launch training step, and a D2H read of this step's result.
        Returns (mean cross-entropy loss of the batch, correct count)."""
        B = x_host.shape[0]
        stage = getattr(self, "_host_stage", None)
        if stage is None or stage[0].shape != x_host.shape:
            stage = (torch.empty(x_host.shape, dtype=torch.float32, device=self.device),
                     torch.empty(B, dtype=torch.int64, device=self.device))
            self._host_stage = stage
        if self._host_metrics is None:
            self._host_metrics = torch.zeros(2, dtype=torch.float32).pin_memory()
        stage[0].copy_(x_host, non_blocking=True)
        stage[1].copy_(y_host.reshape(-1), non_blocking=True)
        m = self.train_step(stage[0], stage[1])
        self._host_metrics.copy_(m, non_blocking=True)
        torch.cuda.current_stream(self.device).synchronize()
        return float(self._host_metrics[0]) / B, int(round(float(self._host_metrics[1])))

    def train_steps_host(self, batches):
        """Pipelined end-to-end training over an iterable of HOST batches ``(x, y)`` (pinned memory for
        truly asynchronous DMA).  Generator: yields ``(mean loss, correct count)`` for every step, in
        order.  Each step still copies ITS inputs host->device and has ITS result read back
        device->host, but the copy of step s+1 runs on a side stream while step s computes, and the
        result of step s is read while step s+1 is already queued (one-step lag), so neither PCIe
        transfer sits on the critical path."""
        dev = self.device
        main = torch.cuda.current_stream(dev)
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(dev)
        cs = self._copy_stream
        slots = [None, None]          # double-buffered device batches
        copied = [torch.cuda.Event(), torch.cuda.Event()]
        consumed = [torch.cuda.Event(), torch.cuda.Event()]
        results = [torch.zeros(2, dtype=torch.float32).pin_memory() for _ in range(2)]
        res_ready = [torch.cuda.Event(), torch.cuda.Event()]

        def stage(i, x, y):
            if slots[i] is None or slots[i][0].shape != x.shape:
                slots[i] = (torch.empty(x.shape, dtype=torch.float32, device=dev),
                            torch.empty(y.numel(), dtype=torch.int64, device=dev))
            with torch.cuda.stream(cs):
                cs.wait_event(consumed[i])            # the step that last used this slot has finished
                slots[i][0].copy_(x, non_blocking=True)
                slots[i][1].copy_(y.reshape(-1), non_blocking=True)
                copied[i].record(cs)

        it = iter(batches)
        nxt = next(it, None)
        if nxt is None:
            return
        for e in consumed:
            e.record(main)
        stage(0, *nxt)
        s = 0
        pending = None                                # (slot, batch size) of the step whose result is in flight
        while nxt is not None:
            i = s & 1
            cur_B = nxt[0].shape[0]
            nxt = next(it, None)
            if nxt is not None:
                stage(i ^ 1, *nxt)                    # prefetch the next batch while this one computes
            main.wait_event(copied[i])
            m = self.train_step(slots[i][0], slots[i][1])
            consumed[i].record(main)
            results[i].copy_(m, non_blocking=True)
            res_ready[i].record(main)
            if pending is not None:
                j, Bj = pending
                res_ready[j].synchronize()
                yield float(results[j][0]) / Bj, int(round(float(results[j][1])))
            pending = (i, cur_B)
            s += 1
        j, Bj = pending
        res_ready[j].synchronize()
        yield float(results[j][0]) / Bj, int(round(float(results[j][1])))


def _train_steps_dataset(self, feats_pinned: torch.Tensor, labels_pinned: torch.Tensor, index_batches):
    """See HostStepsMixin.train_steps_dataset."""
    dev, lib = self.device, _native.lib()
    if not (feats_pinned.is_pinned() and labels_pinned.is_pinned()):
        raise ValueError("train_steps_dataset: the dataset tensors must live in pinned host memory (tensor.pin_memory())")
    if feats_pinned.dtype != torch.float32 or labels_pinned.dtype != torch.int64:
        raise ValueError("train_steps_dataset: features fp32, labels int64")
    T, I = feats_pinned.shape[1], feats_pinned.shape[2]
    labels_flat = labels_pinned.view(-1)
    main = torch.cuda.current_stream(dev)
    if self._copy_stream is None:
        self._copy_stream = torch.cuda.Stream(dev)
    cs = self._copy_stream
    slots, idx_pin, idx_dev = [None, None], [None, None], [None, None]
    copied = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]
    results = [torch.zeros(2, dtype=torch.float32).pin_memory() for _ in range(2)]
    res_ready = [torch.cuda.Event(), torch.cuda.Event()]

    def stage(i, idx):
        B = idx.numel()
        if slots[i] is None or slots[i][0].shape[0] != B:
            slots[i] = (torch.empty(B, T, I, dtype=torch.float32, device=dev), torch.empty(B, dtype=torch.int64, device=dev))
            idx_pin[i] = torch.empty(B, dtype=torch.int64).pin_memory()
            idx_dev[i] = torch.empty(B, dtype=torch.int64, device=dev)
        with torch.cuda.stream(cs):
            cs.wait_event(consumed[i])            # the step that last used this slot (and its index buffer) has finished
            idx_pin[i].copy_(idx)
            idx_dev[i].copy_(idx_pin[i], non_blocking=True)
            # the GPU pulls the batch's rows out of the pinned dataset itself (peer reads over PCIe)
            _native.check(lib.b200rnn_gather_batch_capped(feats_pinned.data_ptr(), labels_flat.data_ptr(), idx_dev[i].data_ptr(), B,
                                                          T * I, slots[i][0].data_ptr(), slots[i][1].data_ptr(), 64,
                                                          cs.cuda_stream), "gather_batch")   # PCIe-bound: 64 CTAs saturate it
            copied[i].record(cs)

    it = iter(index_batches)
    nxt = next(it, None)
    if nxt is None:
        return
    for e in consumed:
        e.record(main)
    stage(0, nxt)
    s, pending = 0, None
    while nxt is not None:
        i = s & 1
        cur_B = nxt.numel()
        nxt = next(it, None)
        if nxt is not None:
            stage(i ^ 1, nxt)
        main.wait_event(copied[i])
        m = self.train_step(slots[i][0], slots[i][1])
        consumed[i].record(main)
        results[i].copy_(m, non_blocking=True)
        res_ready[i].record(main)
        if pending is not None:
            j, Bj = pending
            res_ready[j].synchronize()
            yield float(results[j][0]) / Bj, int(round(float(results[j][1])))
        pending = (i, cur_B)
        s += 1
    j, Bj = pending
    res_ready[j].synchronize()
    yield float(results[j][0]) / Bj, int(round(float(results[j][1])))


class FusedTrainEngine(HostStepsMixin):
    METRIC_RING = 1024

    def __init__(self, model: MotionModel, optimizer: Optional[_FlatOptimizer] = None, lr: float = 0.0025,
                 arena: Optional[ParamArena] = None, comm=None):
        p0 = next(model.parameters())
        if not p0.is_cuda:
            raise ValueError("FusedTrainEngine needs the model on a CUDA device")
        self.lib = _native.lib()
        # forward engine: "ffma" = CUDA-core persistent kernel (default: fastest at H=32, where the
        # recurrence is bound by the serial chain and a tcgen05.mma costs ~64 cycles to issue
        # whatever its N — see DESIGN.md), "tc" = tcgen05/TMEM recurrent kernel (3xTF32)
        self.fwd_kind = os.environ.get("B200RNN_FWD", "ffma")
        self.model = model
        lib = self.lib
        self._fwd_ex = None   # forward that reads the batch through an index slice and takes hand-over flags
        if model.cell == "gru":   # GRU members of the fused family (no tcgen05 variant)
            self._fwd = lib.b200rnn_gru_fwd
            self._fwd_ex = lib.b200rnn_gru_fwd_ex
            self._bwd_ex = lambda d, p, x, idx, st, y, scale, scratch, fl, sp: lib.b200rnn_gru_bwd_ex(
                d, p, x, idx, st, y, scale, None, None, scratch, fl, sp)
        else:
            self._fwd = lib.b200rnn_lstm_fwd_tc if self.fwd_kind == "tc" else lib.b200rnn_lstm_fwd
            if self.fwd_kind != "tc":
                self._fwd_ex = lib.b200rnn_lstm_fwd_ex
            self._bwd_ex = lambda d, p, x, idx, st, y, scale, scratch, fl, sp: lib.b200rnn_lstm_bwd_ex(
                d, p, x, idx, st, y, scale, None, None, None, scratch, fl, sp)
        self.device = p0.device
        self.comm = comm
        self.arena = arena if arena is not None else (optimizer.arena if optimizer is not None else ParamArena(model))
        self.optimizer = optimizer if optimizer is not None else FusedAdam(self.arena, lr=lr)
        self.world = 1 if comm is None else getattr(comm, "grad_world", comm.world)
        self._bufs: Dict[Tuple[int, int], dict] = {}
        self.metrics = torch.zeros(self.METRIC_RING, 2, dtype=torch.float32, device=self.device)
        self.eval_metrics = torch.zeros(256, dtype=torch.float32, device=self.device)
        self.n_steps = 0
        self._host_metrics = None
        self._copy_stream = None
        d = self._desc(1, 1)
        if not self.lib.b200rnn_rnn_small_supported(C.byref(d)):
            raise NotImplementedError(
                f"no fused kernel family covers cell={model.cell} I={model.input_dim} H={model.hidden_dim} "
                f"L={model.layer_dim} C={model.output_dim}")

    # -- descriptor with offsets taken from the arena --
    def _desc(self, B: int, T: int) -> ModelDesc:
        m, a = self.model, self.arena
        d = ModelDesc()
        d.cell = CELL_ID[m.cell]
        d.input_size, d.hidden_size, d.num_layers = m.input_dim, m.hidden_dim, m.layer_dim
        d.batch, d.seq_len, d.num_classes = B, T, m.output_dim
        for l in range(m.layer_dim):
            d.off_wih[l] = a.offset(f"lstm.weight_ih_l{l}")
            d.off_whh[l] = a.offset(f"lstm.weight_hh_l{l}")
            d.off_bih[l] = a.offset(f"lstm.bias_ih_l{l}")
            d.off_bhh[l] = a.offset(f"lstm.bias_hh_l{l}")
        d.off_fcw = a.offset("fc.weight")
        d.off_fcb = a.offset("fc.bias")
        d.n_params = a.numel
        return d

    def _fwd_wave(self, d: ModelDesc) -> int:
        """1 if the forward of this shape is a layer wavefront (the tcgen05 forward never is)."""
        if self.fwd_kind == "tc" and self.model.cell == "lstm":
            return 0
        return int(self.lib.b200rnn_rnn_small_fwd_wave(C.byref(d)))

    def _get(self, B: int, T: int) -> dict:
        key = (B, T)
        buf = self._bufs.get(key)
        if buf is None:
            d = self._desc(B, T)
            rb, nc = C.c_int32(), C.c_int32()
            _native.check(self.lib.b200rnn_rnn_small_tiling(C.byref(d), C.byref(rb), C.byref(nc)), "tiling")
            buf = dict(
                desc=d, n_cta=nc.value, rows_per_cta=rb.value,
                # kernels per step: forward, backward, reduce+update; a wavefront kernel whose buffers do not hold
                # the sentinel yet (first use, or after an evaluation) is preceded by one fill launch
                fwd_wave=bool(self._fwd_wave(d)), bwd_wave=bool(self.lib.b200rnn_rnn_small_bwd_wave(C.byref(d))),
                h_armed=False, dx_armed=False,
                stash=torch.empty(self.lib.b200rnn_rnn_stash_floats(C.byref(d)), dtype=torch.float32, device=self.device),
                scratch=torch.empty(self.lib.b200rnn_rnn_bwd_scratch_floats(C.byref(d)), dtype=torch.float32, device=self.device),
                x=None, y=None)
            self._bufs[key] = buf
        return buf

    # -- one optimisation step on a device batch --
    ARMED, REARM = 1, 2   # B200RNN_STEP_* (include/b200rnn.h)

    def train_step(self, x: torch.Tensor, y: torch.Tensor, idx: Optional[torch.Tensor] = None) -> torch.Tensor:
        """x [B,T,I] fp32, y [B] or [B,1] int64, both on the engine's device.  With ``idx`` ([B] int64 on the
        device) x / y are the whole device-resident dataset and batch row b is row ``idx[b]`` of them.
        Returns a device view [2] = (sum of per-sample CE losses, correct count) for this batch."""
        T = x.shape[1]
        B = x.shape[0] if idx is None else idx.numel()
        if x.dtype != torch.float32 or not x.is_contiguous() or x.device != self.device:
            raise ValueError("train_step: x must be a contiguous fp32 tensor on the engine's device")
        y = y.reshape(-1)
        if y.dtype != torch.int64 or not y.is_contiguous() or y.device != self.device:
            raise ValueError("train_step: labels must be a contiguous int64 tensor on the engine's device")
        if idx is not None and (idx.dtype != torch.int64 or not idx.is_contiguous() or idx.device != self.device):
            raise ValueError("train_step: idx must be a contiguous int64 tensor on the engine's device")
        buf = self._get(B, T)
        d = buf["desc"]
        sp = _native.stream_ptr(self.device)
        a, opt = self.arena, self.optimizer
        n_launch = 3
        if self._fwd_ex is None:   # tcgen05 forward: plain entry point, explicit gather in front
            if idx is not None:
                x, y = self._gather(buf, x, y, idx)
                idx = None
                n_launch += 1
            _native.check(self._fwd(C.byref(d), a.flat.data_ptr(), x.data_ptr(), buf["stash"].data_ptr(), 1, sp), "lstm_fwd")
            buf["h_armed"] = False
        else:
            f_fl = self.ARMED if buf["h_armed"] else 0
            n_launch += int(buf["fwd_wave"] and not buf["h_armed"])
            _native.check(self._fwd_ex(C.byref(d), a.flat.data_ptr(), x.data_ptr(), _native.ptr(idx),
                                       buf["stash"].data_ptr(), 1, f_fl, sp), "rnn_fwd")
            buf["h_armed"] = False
        # the consumers put the wavefront sentinels back, so the next step needs no fill launch
        rearm = self._fwd_ex is not None and (buf["fwd_wave"] or buf["bwd_wave"])
        b_fl = (self.ARMED if buf["dx_armed"] else 0) | (self.REARM if rearm else 0)
        n_launch += int(buf["bwd_wave"] and not buf["dx_armed"])
        _native.check(self._bwd_ex(C.byref(d), a.flat.data_ptr(), x.data_ptr(), _native.ptr(idx), buf["stash"].data_ptr(),
                                   y.data_ptr(), 1.0 / B, buf["scratch"].data_ptr(), b_fl, sp), "rnn_bwd")
        buf["h_armed"], buf["dx_armed"] = rearm, rearm and buf["bwd_wave"]
        opt.step_count += 1
        slot = self.metrics[self.n_steps % self.METRIC_RING]
        o = opt.desc(grad_scale=1.0 / self.world)
        if self.comm is None or (self.world == 1 and not getattr(self.comm, "always", False)):
            _native.check(self.lib.b200rnn_fused_update(
                buf["scratch"].data_ptr(), buf["n_cta"], a.numel, a.flat.data_ptr(), opt.exp_avg.data_ptr(),
                opt.exp_avg_sq.data_ptr(), C.byref(o), None, slot.data_ptr(), sp), "fused_update")
        else:
            self.comm.allreduce_update(buf["scratch"], buf["n_cta"], a, opt, o, slot)
        _native.count_launch(n_launch)
        self.n_steps += 1
        return slot

    def _gather(self, buf: dict, feats: torch.Tensor, labels: torch.Tensor, idx: torch.Tensor):
        B, T, I = idx.numel(), feats.shape[1], feats.shape[2]
        if buf["x"] is None:
            buf["x"] = torch.empty(B, T, I, dtype=torch.float32, device=self.device)
            buf["y"] = torch.empty(B, dtype=torch.int64, device=self.device)
        _native.check(self.lib.b200rnn_gather_batch(feats.data_ptr(), labels.data_ptr(), idx.data_ptr(), B, T * I,
                                                    buf["x"].data_ptr(), buf["y"].data_ptr(),
                                                    _native.stream_ptr(self.device)), "gather_batch")
        return buf["x"], buf["y"]

    def train_step_gather(self, feats: torch.Tensor, labels: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
        """Train on rows ``idx`` of a device-resident dataset (feats [N,T,I], labels [N] or [N,1]): the gather of
        the reference's DataLoader (src/motion/trainer/base.py:107) happens inside the forward / backward loaders."""
        return self.train_step(feats, labels, idx)

    @torch.no_grad()
    def evaluate(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """Forward + head only. Returns device view [2] = (loss sum, correct count)."""
        B, T, _ = x.shape
        buf = self._get(B, T)
        d = buf["desc"]
        y = y.reshape(-1)
        sp = _native.stream_ptr(self.device)
        a, m = self.arena, self.model
        H, L = m.hidden_dim, m.layer_dim
        _native.check(self._fwd(C.byref(d), a.flat.data_ptr(), x.data_ptr(), buf["stash"].data_ptr(), 0, sp), "lstm_fwd")
        buf["h_armed"] = False   # the stash now holds activations, not sentinels
        h_last = buf["stash"][((L - 1) * T + (T - 1)) * B * H:][:B * H]
        _native.check(self.lib.b200rnn_head_fwd(h_last.data_ptr(), H, m.fc.weight.data_ptr(), m.fc.bias.data_ptr(),
                                                y.data_ptr(), B, H, m.output_dim, None, self.eval_metrics.data_ptr(), sp), "head_fwd")
        _native.count_launch(3 + int(buf["fwd_wave"] and self._fwd_ex is not None))
        return self.eval_metrics[:2]
