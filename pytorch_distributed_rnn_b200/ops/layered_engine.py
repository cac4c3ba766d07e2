"""LayeredTrainEngine — the training step of a LARGE-hidden motion classifier (BASELINE.json config 4:
2 x 1024, T = 512, bf16 operands) with per-layer gradient buckets that are all-reduced while the
backward pass of the layers below is still running.

The reference gets this behaviour from ``DistributedDataParallel(model)``: c10d's Reducer hooks fire a
bucket's all-reduce during ``loss.backward()`` (src/motion/trainer/ddp.py:19, trainer/base.py:116), and
the optimizer loop runs afterwards (base.py:117).  Here the step is driven natively, layer by layer:

    main stream                                   communication stream
    -----------                                   --------------------
    K11 gather + transpose (time-major batch)
    forward, all layers (tcgen05 GEMMs + persistent recurrence, csrc/rnn/rnn_large.cu, rnn_persist.cu)
    K5 head + CE + their backward (csrc/head)  ->  wait(bucket fc)   all-reduce + Adam on the fc slice
    backward of layer L-1 ... dW GEMM tiles    ->  wait(bucket L-1)  all-reduce + Adam on layer L-1's slice
    backward of layer L-2 (runs meanwhile)     ->  wait(bucket L-2)  ...
    ...
    join

``wait(bucket)`` is not a stream event: it is the reducer's one-thread kernel spinning on a device counter
that the weight-gradient GEMM tiles (and the bias-sum / head blocks) increment as they retire
(csrc/runtime/reducer.cpp, csrc/gemm/gemm_tc.cu) — the collective starts the moment the bucket's last tile
is out.  The collective itself is the in-kernel NVLink all-reduce fused with the gradient average and the
Adam step (two-shot / NVLS for the multi-megabyte layer buckets, LL one-shot for the head), picked per
bucket size from the measured table (:mod:`..parallel.symm`).  On one GPU the same pipeline runs with the
fused optimizer in place of the collective, so Adam on layer l overlaps the backward of layer l-1.

``timeline=True`` stamps %globaltimer at bucket trigger / collective end / layer-backward begin and end;
``last_timeline()`` returns them in microseconds relative to the step's first stamp.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, List, Optional, Tuple

import torch

from .. import _native
from .._native import ModelDesc
from ..models.motion import MotionModel
from ..utils.arena import ParamArena
from .engine import HostStepsMixin
from .optim import FusedAdam, _FlatOptimizer
from .rnn_function import CELL_ID


class LayeredTrainEngine(HostStepsMixin):
    METRIC_RING = 1024

    def __init__(self, model: MotionModel, optimizer: Optional[_FlatOptimizer] = None, lr: float = 0.0025,
                 arena: Optional[ParamArena] = None, comm=None, timeline: bool = False):
        p0 = next(model.parameters())
        if not p0.is_cuda:
            raise ValueError("LayeredTrainEngine needs the model on a CUDA device")
        if model.hidden_dim % 8 != 0 or model.output_dim > 32:
            raise NotImplementedError(f"large-hidden path needs hidden % 8 == 0 and <= 32 classes "
                                      f"(H={model.hidden_dim}, C={model.output_dim})")
        self.lib = _native.lib()
        self.model, self.device, self.comm = model, p0.device, comm
        self.arena = arena if arena is not None else (optimizer.arena if optimizer is not None else ParamArena(model))
        self.optimizer = optimizer if optimizer is not None else FusedAdam(self.arena, lr=lr)
        self.world = 1 if comm is None else getattr(comm, "grad_world", comm.world)
        self.metrics = torch.zeros(self.METRIC_RING, 2, dtype=torch.float32, device=self.device)
        self.eval_metrics = torch.zeros(256, dtype=torch.float32, device=self.device)
        self.n_steps = 0
        self.timeline = timeline
        self._bufs: Dict[Tuple[int, int], dict] = {}
        self._copy_stream = None
        self._host_metrics = None
        self.comm_stream = torch.cuda.Stream(self.device)
        # ---- gradient buckets: the head, then one per recurrent layer from the top down (the order in
        # which their gradients complete); each is a contiguous slice of the arena ----
        a, L = self.arena, model.layer_dim
        self.buckets: List[Tuple[str, int, int]] = [("fc", a.offset("fc.weight"), a.numel)]
        for l in range(L - 1, -1, -1):
            lo = a.offset(f"lstm.weight_ih_l{l}")
            o, k, _ = a.index[f"lstm.bias_hh_l{l}"]
            self.buckets.append((f"layer{l}", lo, o + k))
        lo_arr = (C.c_int64 * len(self.buckets))(*[b[1] for b in self.buckets])
        hi_arr = (C.c_int64 * len(self.buckets))(*[b[2] for b in self.buckets])
        self._reducer = C.c_void_p()
        _native.check(self.lib.b200rnn_reducer_create(len(self.buckets), lo_arr, hi_arr, C.byref(self._reducer)), "reducer_create")

    def __del__(self):
        try:
            if getattr(self, "_reducer", None):
                self.lib.b200rnn_reducer_destroy(self._reducer)
                self._reducer = None
        except Exception:  # noqa: BLE001 - interpreter shutdown
            pass

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

    def _get(self, B: int, T: int) -> dict:
        key = (B, T)
        buf = self._bufs.get(key)
        if buf is None:
            d = self._desc(B, T)
            m, dev, lib = self.model, self.device, self.lib
            H, L = m.hidden_dim, m.layer_dim
            buf = dict(
                desc=d,
                stash=torch.empty(lib.b200rnn_rnn_large_stash_bytes(C.byref(d)), dtype=torch.uint8, device=dev),
                scratch=torch.empty(lib.b200rnn_rnn_large_scratch_bytes(C.byref(d)), dtype=torch.uint8, device=dev),
                x_tm=torch.empty(T, B, m.input_dim, dtype=torch.float32, device=dev),
                d_hn=torch.zeros(L, B, H, dtype=torch.float32, device=dev),      # only the top layer's slice is ever written
                head=torch.empty(lib.b200rnn_head_ce_scratch_floats(B, m.output_dim), dtype=torch.float32, device=dev),
                h_last_off=lib.b200rnn_rnn_large_h_offset(C.byref(d), L - 1) + (T - 1) * B * H * 4,
            )
            self._bufs[key] = buf
        return buf

    # -- what the communication stream does for one bucket once its trigger has fired --
    def _bucket_update(self, lo: int, hi: int, optdesc) -> None:
        a, opt, n = self.arena, self.optimizer, hi - lo
        sp = _native.stream_ptr(self.device)
        g, p = a.grad.data_ptr() + 4 * lo, a.flat.data_ptr() + 4 * lo
        m, v = opt.exp_avg.data_ptr() + 4 * lo, opt.exp_avg_sq.data_ptr() + 4 * lo
        comm = self.comm
        if comm is None or self.world == 1:
            _native.check(self.lib.b200rnn_fused_update(g, 1, n, p, m, v, C.byref(optdesc), None, None, sp), "fused_update")
        elif hasattr(comm, "allreduce_update_slice"):
            comm.allreduce_update_slice(a, opt, optdesc, lo, hi)
        else:   # NCCL baseline: all-reduce the slice, then the fused optimizer (grad_scale = 1 / world)
            import torch.distributed as dist
            dist.all_reduce(a.grad[lo:hi])
            _native.check(self.lib.b200rnn_fused_update(g, 1, n, p, m, v, C.byref(optdesc), None, None, sp), "fused_update")

    def train_step(self, x: torch.Tensor, y: torch.Tensor, idx: Optional[torch.Tensor] = None) -> torch.Tensor:
        """x [B,T,I] fp32 and y [B] / [B,1] int64 on the engine's device — or, with ``idx`` ([B] int64), the whole
        device-resident dataset of which batch row b is row ``idx[b]``.  Returns a device view [2] =
        (sum of per-sample CE losses, correct count)."""
        T = x.shape[1]
        B = x.shape[0] if idx is None else idx.numel()
        if x.dtype != torch.float32 or not x.is_contiguous() or x.device != self.device:
            raise ValueError("train_step: x must be a contiguous fp32 tensor on the engine's device")
        y = y.reshape(-1)
        if y.dtype != torch.int64 or not y.is_contiguous():
            raise ValueError("train_step: labels must be a contiguous int64 tensor")
        lib, m, a, opt = self.lib, self.model, self.arena, self.optimizer
        buf = self._get(B, T)
        d = buf["desc"]
        H, L, Cn = m.hidden_dim, m.layer_dim, m.output_dim
        main = torch.cuda.current_stream(self.device)
        cs = self.comm_stream
        sp, csp = main.cuda_stream, cs.cuda_stream
        red = self._reducer
        cs.wait_stream(main)   # step boundary only: earlier main-stream work (broadcasts, evaluation) is done
        _native.check(lib.b200rnn_gather_batch_tm(x.data_ptr(), _native.ptr(idx), B, T, m.input_dim, buf["x_tm"].data_ptr(), sp), "gather_tm")
        _native.check(lib.b200rnn_rnn_large_fwd(C.byref(d), a.flat.data_ptr(), buf["x_tm"].data_ptr(), buf["stash"].data_ptr(),
                                                buf["scratch"].data_ptr(), sp), "rnn_large_fwd")
        opt.step_count += 1
        o = opt.desc(grad_scale=1.0 / self.world)
        slot = self.metrics[self.n_steps % self.METRIC_RING]
        cnt = C.c_int32(0)
        # ---- head + CE + their gradients; bucket 0 ----
        h_last = buf["stash"].data_ptr() + buf["h_last_off"]
        _native.check(lib.b200rnn_head_ce(
            h_last, H, a.flat.data_ptr() + 4 * d.off_fcw, a.flat.data_ptr() + 4 * d.off_fcb, y.data_ptr(), _native.ptr(idx),
            B, H, Cn, 1.0 / B, buf["d_hn"][L - 1].data_ptr(), H, a.grad.data_ptr() + 4 * d.off_fcw,
            a.grad.data_ptr() + 4 * d.off_fcb, slot.data_ptr(), buf["head"].data_ptr(),
            lib.b200rnn_reducer_counter(red, 0), C.byref(cnt), sp), "head_ce")
        self._consume(0, cnt.value, o)
        # ---- recurrent layers, top down; bucket 1 + (L-1-l) ----
        for l in range(L - 1, -1, -1):
            b = 1 + (L - 1 - l)
            if self.timeline:
                lib.b200rnn_reducer_stamp(red, b, 2, sp)
            _native.check(lib.b200rnn_rnn_large_bwd_layer(
                C.byref(d), a.flat.data_ptr(), buf["x_tm"].data_ptr(), buf["stash"].data_ptr(), buf["scratch"].data_ptr(), l,
                None, buf["d_hn"].data_ptr(), None, a.grad.data_ptr(), None,
                lib.b200rnn_reducer_counter(red, b), C.byref(cnt), sp), "rnn_large_bwd_layer")
            if self.timeline:
                lib.b200rnn_reducer_stamp(red, b, 3, sp)
            self._consume(b, cnt.value, o)
        main.wait_stream(cs)   # the next forward reads the updated parameters
        # kernels of this repo per step: gather, per layer 6 forward (3 casts, bias add, projection GEMM, recurrence)
        # and 11 backward (2 weight transposes, recurrence, dX GEMM, 3 operand transposes, 2 dW GEMMs, 2 bias sums),
        # the head's 2, and per bucket the trigger wait + the collective / optimizer
        _native.count_launch(1 + 17 * L + 2 + 2 * len(self.buckets))
        self.n_steps += 1
        return slot

    def _consume(self, bucket: int, counts: int, optdesc) -> None:
        """Communication stream: wait (on the device) for the bucket's producers, then all-reduce + update."""
        lib, red = self.lib, self._reducer
        _native.check(lib.b200rnn_reducer_expect(red, bucket, counts), "reducer_expect")
        _, lo, hi = self.buckets[bucket]
        with torch.cuda.stream(self.comm_stream):
            _native.check(lib.b200rnn_reducer_wait(red, bucket, self.comm_stream.cuda_stream), "reducer_wait")
            self._bucket_update(lo, hi, optdesc)
            if self.timeline:
                lib.b200rnn_reducer_stamp(red, bucket, 1, self.comm_stream.cuda_stream)

    def last_timeline(self) -> dict:
        """Stamps of the most recent step (synchronises): per bucket, microseconds since the step's first stamp."""
        n = len(self.buckets)
        raw = (C.c_uint64 * (4 * n))()
        st = C.c_int32(0)
        _native.check(self.lib.b200rnn_reducer_read(self._reducer, raw, C.byref(st)), "reducer_read")
        vals = [int(v) for v in raw]
        t0 = min(v for v in vals if v > 0) if any(vals) else 0
        us = lambda v: None if v == 0 else (v - t0) / 1e3
        out = {"status": st.value, "buckets": {}}
        for i, (name, lo, hi) in enumerate(self.buckets):
            trig, done, bwd0, bwd1 = vals[4 * i:4 * i + 4]
            out["buckets"][name] = {"bytes": 4 * (hi - lo), "trigger_us": us(trig), "collective_done_us": us(done),
                                    "backward_begin_us": us(bwd0), "backward_end_us": us(bwd1)}
        return out

    def reducer_status(self) -> int:
        st = C.c_int32(0)
        _native.check(self.lib.b200rnn_reducer_read(self._reducer, None, C.byref(st)), "reducer_read")
        return st.value

    def train_step_gather(self, feats: torch.Tensor, labels: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
        return self.train_step(feats, labels, idx)

    @torch.no_grad()
    def evaluate(self, x: torch.Tensor, y: torch.Tensor) -> torch.Tensor:
        """Forward + head only.  Returns a device view [2] = (loss sum, correct count)."""
        B, T, _ = x.shape
        lib, m, a = self.lib, self.model, self.arena
        buf = self._get(B, T)
        d = buf["desc"]
        y = y.reshape(-1)
        sp = _native.stream_ptr(self.device)
        _native.check(lib.b200rnn_gather_batch_tm(x.data_ptr(), None, B, T, m.input_dim, buf["x_tm"].data_ptr(), sp), "gather_tm")
        _native.check(lib.b200rnn_rnn_large_fwd(C.byref(d), a.flat.data_ptr(), buf["x_tm"].data_ptr(), buf["stash"].data_ptr(),
                                                buf["scratch"].data_ptr(), sp), "rnn_large_fwd")
        h_last = buf["stash"].data_ptr() + buf["h_last_off"]
        _native.check(lib.b200rnn_head_fwd(h_last, m.hidden_dim, m.fc.weight.data_ptr(), m.fc.bias.data_ptr(), y.data_ptr(), B,
                                           m.hidden_dim, m.output_dim, None, self.eval_metrics.data_ptr(), sp), "head_fwd")
        return self.eval_metrics[:2]
