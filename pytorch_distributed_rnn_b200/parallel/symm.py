"""SymmComm — the product gradient path: symmetric memory + in-kernel NVLink collectives.

Each rank owns one symmetric region (flags | inbox | heap) allocated by the native runtime and
mapped into every other rank with CUDA IPC, so kernels address peers directly over NVLink /
NVSwitch.  On top of it:

* ``allreduce_update`` (K7): ONE kernel per step reduces the per-CTA gradient rows of the
  backward kernel, all-reduces them (one-shot for small vectors such as the reference's 56.6 KB
  gradient, two-shot reduce-scatter + parameter all-gather for large ones), averages, and
  applies Adam/SGD — replacing c10d's Reducer + MPI_Allreduce + the optimizer loop
  (src/motion/trainer/ddp.py:19, trainer/base.py:117) and Horovod's fused all-reduce
  (trainer/horovod.py:35);
* ``broadcast_flat`` (K8), ``put`` / ``get`` (K9), ``barrier`` — constructor broadcast of DDP,
  ``hvd.broadcast_parameters``, and the send/recv of src/example/example_distributed.py;
* ``make_arena`` places the flat parameter arena inside the symmetric heap so that the two-shot
  all-gather and the broadcast can store straight into every rank's parameters.

A watchdog in every flag wait turns a dead peer into a status bit instead of a hung GPU
(``check()``), and ``delay_ns`` injects latency before every flag publication — the analogue
of the reference's ``tc netem`` experiment (fabfile.py:130-191).
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional

import torch
import torch.distributed as dist

from .. import _native
from .._native import CommCtx
from ..utils.arena import ParamArena
from .bootstrap import NativeBootstrap, probe_topology
from .process_group import RankInfo, rank_info


class _External:
    """Expose a raw device range to torch (zero-copy) through __cuda_array_interface__."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}


ALGO_AUTO, ALGO_ONESHOT, ALGO_TWOSHOT, ALGO_NVLS, ALGO_LL = 0, 1, 2, 3, 4
ALGO_NAMES = {ALGO_ONESHOT: "oneshot", ALGO_TWOSHOT: "twoshot", ALGO_NVLS: "nvls", ALGO_LL: "ll"}

# Per-world-size crossovers between the all-reduce variants (vector bytes).  The defaults are what the
# round-1 sweeps suggested; `bench/allreduce_sweep.py --write-table` replaces them with the measured
# crossovers of the box it ran on and persists them in profiles/allreduce_table.json, which is loaded here
# (B200RNN_ALGO_TABLE points elsewhere) — SURVEY §5.8 item 3: "picked per bucket from a measured table".
DEFAULT_TABLE = {"ll_max_bytes": 256 << 10, "oneshot_max_bytes": 0, "nvls_min_bytes": 1 << 20}


def load_algo_table(world: int, path: Optional[str] = None) -> dict:
    import json
    from pathlib import Path
    table = dict(DEFAULT_TABLE)
    table["source"] = "defaults"
    p = Path(path or os.environ.get("B200RNN_ALGO_TABLE") or
             Path(__file__).resolve().parents[2] / "profiles" / "allreduce_table.json")
    try:
        doc = json.loads(p.read_text())
        row = doc.get("worlds", {}).get(str(world))
        if row:
            table.update({k: int(v) for k, v in row.items() if k in DEFAULT_TABLE})
            table["source"] = str(p)
    except (OSError, ValueError):
        pass
    return table


class SymmComm:
    kind = "symm"
    MAILBOX_BYTES = 1 << 16  # per source rank, for put/get

    def __init__(self, info: Optional[RankInfo] = None, inbox_floats: Optional[int] = None,
                 heap_bytes: Optional[int] = None, max_blocks: int = 1 << 16, delay_ns: int = 0,
                 boot: Optional[NativeBootstrap] = None):
        self.info = info or rank_info()
        self.rank, self.world, self.device = self.info.rank, self.info.world_size, self.info.device
        if self.device.type != "cuda":
            raise RuntimeError("SymmComm needs CUDA devices")
        self.lib = _native.lib()
        inbox_floats = inbox_floats or int(os.environ.get("B200RNN_INBOX_FLOATS", 4 << 20))
        heap_bytes = heap_bytes or int(os.environ.get("B200RNN_HEAP_BYTES", 256 << 20))
        self.topology = probe_topology()
        # ---- rendezvous: rank 0 opens an ephemeral port and tells the others through the
        # already-initialised torch process group (or B200RNN_BOOT_PORT without one) ----
        if boot is None:
            if self.world > 1:
                if dist.is_initialized():
                    box = [None]
                    fd = -1
                    if self.rank == 0:
                        port, fd = NativeBootstrap.listen(0)
                        box[0] = port
                    dist.broadcast_object_list(box, src=0)
                    boot = NativeBootstrap(self.rank, self.world, os.environ.get("MASTER_ADDR", "127.0.0.1"),
                                           box[0], fd)
                else:
                    boot = NativeBootstrap(self.rank, self.world, os.environ.get("MASTER_ADDR", "127.0.0.1"))
            else:
                boot = NativeBootstrap(0, 1)
        self.boot = boot
        ctx = CommCtx()
        delay_ns = delay_ns or int(os.environ.get("B200RNN_DELAY_NS", "0"))
        ctx.rank, ctx.world, ctx.max_blocks, ctx.delay_ns = self.rank, self.world, max_blocks, delay_ns
        # watchdog of the in-kernel waits: B200RNN_COMM_TIMEOUT_MS (default 60 s, resolved natively when 0 here)
        ctx.inbox_floats, ctx.timeout_cycles, ctx.heap_bytes = inbox_floats, 0, heap_bytes
        ctx.ll_floats = int(os.environ.get("B200RNN_LL_FLOATS", 256 << 10))   # LL one-shot up to 1 MiB vectors
        ctx.loss_ppm = int(os.environ.get("B200RNN_LOSS_PPM", "0"))
        ctx.loss_rto_ns = int(os.environ.get("B200RNN_LOSS_RTO_NS", "0"))     # 0: the native default, 200 us
        region = self.lib.b200rnn_comm_layout(C.byref(ctx))
        if region <= 0:
            raise _native.NativeError("comm_layout failed")
        # Region allocation: the VMM + NVSwitch-multicast form when the node supports NVLS (so that the
        # in-switch reduction is available), else cudaMalloc + CUDA IPC.  B200RNN_SYMM=ipc|vmm overrides.
        self.symm = C.c_void_p()
        self.mc_base = None
        self._vmm = False
        ptrs = (C.c_void_p * _native.MAX_PEERS)()
        mode = os.environ.get("B200RNN_SYMM", "auto")
        want_vmm = mode == "vmm" or (mode == "auto" and self.world > 1 and all(self.topology["multicast_supported"]))
        if want_vmm:
            mc = C.c_void_p()
            rc = self.lib.b200rnn_symm_vmm_create(boot.handle, self.rank, self.world, self.device.index, region,
                                                  int(boot.port) & 0x7FFFFFFF, 1, C.byref(self.symm))
            ok = torch.tensor([1 if rc == 0 else 0], device=self.device)
            if self.world > 1 and dist.is_initialized():
                dist.all_reduce(ok, op=dist.ReduceOp.MIN)     # all ranks take the same path
            if int(ok.item()) == 1:
                _native.check(self.lib.b200rnn_symm_vmm_pointers(self.symm, ptrs, C.byref(mc)), "symm_vmm_pointers")
                self.mc_base = mc.value
                self._vmm = True
            else:
                why = self.lib.b200rnn_last_error().decode("utf-8", "replace") if rc != 0 else "a peer failed"
                if mode == "vmm":
                    raise _native.NativeError(f"symm_vmm_create failed: {why}")
                import logging
                logging.warning("VMM/multicast symmetric memory unavailable (%s); using cudaMalloc + CUDA IPC", why)
                if rc == 0:
                    self.lib.b200rnn_symm_vmm_destroy(self.symm)
                self.symm = C.c_void_p()
        if not self._vmm:
            _native.check(self.lib.b200rnn_symm_create(boot.handle, self.device.index, region, C.byref(self.symm)), "symm_create")
            _native.check(self.lib.b200rnn_symm_pointers(self.symm, ptrs), "symm_pointers")
        for r in range(self.world):
            ctx.base[r] = ptrs[r]
        self.ctx = ctx
        self.region_bytes = region
        self.local_base = ptrs[self.rank]
        self.lib.b200rnn_comm_forget(self.local_base)
        self.heap = torch.as_tensor(_External(self.local_base + ctx.off_heap, heap_bytes), device=self.device)
        self._heap_top = self.world * self.MAILBOX_BYTES  # mailboxes first
        self._stage_bytes = min(16 << 20, heap_bytes // 8)
        self._stage_off = self.heap_alloc(self._stage_bytes)  # reusable staging slot for broadcasts
        self._seq_ar = 0
        self._seq_bc = 0
        self._seq_bar = 0
        self._seq_put = [0] * self.world
        self._seq_get = [0] * self.world
        self.algo = int(os.environ.get("B200RNN_ALGO", "0"))
        self.table = load_algo_table(self.world)
        self.last_algo = None

    # ---- symmetric heap ----
    def heap_alloc(self, nbytes: int, align: int = 256) -> int:
        off = (self._heap_top + align - 1) // align * align
        if off + nbytes > self.ctx.heap_bytes:
            raise MemoryError(f"symmetric heap exhausted ({off + nbytes} > {self.ctx.heap_bytes} bytes)")
        self._heap_top = off + nbytes
        return off

    def heap_tensor(self, numel: int, dtype=torch.float32):
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        off = self.heap_alloc(nbytes)
        return self.heap[off:off + nbytes].view(dtype), off

    def heap_offset_of(self, t: torch.Tensor) -> Optional[int]:
        lo = self.local_base + self.ctx.off_heap
        p = t.data_ptr()
        if lo <= p and p + t.numel() * t.element_size() <= lo + self.ctx.heap_bytes:
            return p - lo
        return None

    def make_arena(self, module) -> ParamArena:
        n = sum(p.numel() for p in module.parameters())
        flat, off = self.heap_tensor(n)
        arena = ParamArena(module, flat=flat)
        arena.heap_off = off
        # gradient slot of the same (4-padded) size at a symmetric offset: operand of the in-switch reduction
        n4 = (n + 3) // 4 * 4
        arena.grad_heap_off = self.heap_alloc(n4 * 4)
        return arena

    # ---- collectives ----
    def _sp(self):
        return _native.stream_ptr(self.device)

    def pick_algo(self, n: int, in_heap: bool, nvls_ok: bool) -> int:
        """Variant for an n-float vector: the forced one (B200RNN_ALGO) or the measured table's."""
        nbytes, t = n * 4, self.table
        ll_fits = (n + 3) // 4 * 4 <= self.ctx.ll_floats
        algo = self.algo
        if algo == ALGO_NVLS and not nvls_ok:
            raise RuntimeError("B200RNN_ALGO=3 (NVLS) needs a multicast-capable symmetric region and a heap-resident arena")
        if algo == ALGO_LL and not ll_fits:
            raise RuntimeError(f"B200RNN_ALGO=4 (LL): {n} floats exceed the LL inbox ({self.ctx.ll_floats})")
        if algo == ALGO_TWOSHOT and not in_heap:
            algo = ALGO_AUTO
        if algo != ALGO_AUTO:
            return algo
        if self.world == 1 or (ll_fits and nbytes <= t["ll_max_bytes"]):
            return ALGO_LL
        if nvls_ok and nbytes >= t["nvls_min_bytes"]:
            return ALGO_NVLS
        if in_heap and nbytes > t["oneshot_max_bytes"]:
            return ALGO_TWOSHOT
        return ALGO_ONESHOT

    def allreduce_update(self, rows: torch.Tensor, n_rows: int, arena, opt, optdesc, metrics_slot: torch.Tensor) -> None:
        self._seq_ar += 1
        heap_off = getattr(arena, "heap_off", None)
        nvls_ok = self.mc_base is not None and heap_off is not None and hasattr(arena, "grad_heap_off")
        algo = self.pick_algo(arena.numel, heap_off is not None, nvls_ok)
        self.last_algo = algo
        if n_rows > 1 and algo != ALGO_LL:
            # only the LL form reduces the per-CTA rows itself; the streaming forms take one row
            _native.check(self.lib.b200rnn_reduce_rows(rows.data_ptr(), n_rows, arena.numel + 2, arena.grad.data_ptr(), 0,
                                                       self._sp()), "reduce_rows")
            _native.count_launch()
            rows, n_rows = arena.grad, 1
        if algo == ALGO_NVLS:
            _native.check(self.lib.b200rnn_comm_allreduce_update_nvls(
                C.byref(self.ctx), self.mc_base, self._seq_ar, rows.data_ptr(), n_rows, arena.numel, arena.flat.data_ptr(),
                heap_off, arena.grad_heap_off, opt.exp_avg.data_ptr(), opt.exp_avg_sq.data_ptr(), C.byref(optdesc),
                metrics_slot.data_ptr(), self._sp()), "comm_allreduce_update_nvls")
            return
        _native.check(self.lib.b200rnn_comm_allreduce_update(
            C.byref(self.ctx), self._seq_ar, rows.data_ptr(), n_rows, arena.numel, arena.flat.data_ptr(),
            heap_off if heap_off is not None else -1, opt.exp_avg.data_ptr(), opt.exp_avg_sq.data_ptr(),
            C.byref(optdesc), metrics_slot.data_ptr(), algo, self._sp()), "comm_allreduce_update")

    def allreduce_update_slice(self, arena, opt, optdesc, lo: int, hi: int) -> None:
        """K7 on one gradient bucket: all-reduce ``arena.grad[lo:hi]`` (a single row), average, and run the
        optimizer on ``arena.flat[lo:hi]`` and its state — on the CURRENT stream (the layered engine calls this on
        its communication stream, behind the reducer's device-side trigger).  All ranks must call it for the same
        buckets in the same order."""
        self._seq_ar += 1
        n = hi - lo
        heap_off = getattr(arena, "heap_off", None)
        in_heap = heap_off is not None and (heap_off + 4 * lo) % 16 == 0
        nvls_ok = (self.mc_base is not None and in_heap and hasattr(arena, "grad_heap_off")
                   and (arena.grad_heap_off + 4 * lo) % 16 == 0)
        algo = self.pick_algo(n, in_heap, nvls_ok)
        self.last_algo = algo
        g = arena.grad.data_ptr() + 4 * lo
        p, m, v = arena.flat.data_ptr() + 4 * lo, opt.exp_avg.data_ptr() + 4 * lo, opt.exp_avg_sq.data_ptr() + 4 * lo
        if algo == ALGO_NVLS:
            _native.check(self.lib.b200rnn_comm_allreduce_update_nvls(
                C.byref(self.ctx), self.mc_base, self._seq_ar, g, 1, n, p, heap_off + 4 * lo, arena.grad_heap_off + 4 * lo,
                m, v, C.byref(optdesc), None, self._sp()), "comm_allreduce_update_nvls")
            return
        _native.check(self.lib.b200rnn_comm_allreduce_update(
            C.byref(self.ctx), self._seq_ar, g, 1, n, p, heap_off + 4 * lo if in_heap else -1, m, v, C.byref(optdesc), None,
            algo, self._sp()), "comm_allreduce_update")

    def allreduce_flat(self, t: torch.Tensor, average: bool = True, algo: int = ALGO_AUTO) -> torch.Tensor:
        assert t.dtype == torch.float32 and t.is_contiguous()
        n, cap = t.numel(), self.ctx.inbox_floats
        scale = 1.0 / self.world if average else 1.0
        for lo in range(0, n, cap):
            k = min(cap, n - lo)
            self._seq_ar += 1
            a = algo
            if a == ALGO_AUTO:
                ll_ok = (k + 3) // 4 * 4 <= self.ctx.ll_floats and k * 4 <= self.table["ll_max_bytes"]
                a = ALGO_LL if ll_ok else ALGO_ONESHOT
            _native.check(self.lib.b200rnn_comm_allreduce_algo(C.byref(self.ctx), self._seq_ar, t.data_ptr() + 4 * lo, k,
                                                               scale, a, self._sp()), "comm_allreduce")
        _native.count_launch((n + cap - 1) // cap)
        return t

    def broadcast_flat(self, t: torch.Tensor, root: int = 0) -> torch.Tensor:
        if self.world == 1:
            return t
        # every rank must be done writing the destination (e.g. its own initial values) before the root's
        # data may land there, and must have consumed it before the next broadcast: barrier on both sides
        self.barrier()
        nbytes = t.numel() * t.element_size()
        off = self.heap_offset_of(t)
        staged = None
        if off is None or off % 16 != 0:
            # symmetric offsets must match on every rank: all ranks take the same branch for the same call
            staged, off = self.heap, (self._stage_off if nbytes <= self._stage_bytes else self.heap_alloc(nbytes))
            if self.rank == root:   # only the root fills its staging slot: a peer writing its own slot could
                #                     overwrite data the root has already delivered there
                self.heap[off:off + nbytes].copy_(t.contiguous().view(-1).view(torch.uint8))
        self._seq_bc += 1
        _native.check(self.lib.b200rnn_comm_broadcast(C.byref(self.ctx), self._seq_bc, off, nbytes, root, self._sp()),
                      "comm_broadcast")
        if staged is not None:
            t.view(-1).view(torch.uint8).copy_(self.heap[off:off + nbytes])
        # the next broadcast may reuse peers' memory only after everyone has consumed this one
        self.barrier()
        return t

    def barrier(self) -> None:
        self._seq_bar += 1
        _native.check(self.lib.b200rnn_comm_barrier(C.byref(self.ctx), self._seq_bar, self._sp()), "comm_barrier")

    def put(self, t: torch.Tensor, dst: int) -> None:
        """K9: one-sided store of ``t`` into rank ``dst``'s mailbox + release flag (no receiver action)."""
        nbytes = t.numel() * t.element_size()
        if nbytes > self.MAILBOX_BYTES:
            raise ValueError("put: message larger than the mailbox")
        self._seq_put[dst] += 1
        _native.check(self.lib.b200rnn_comm_put(C.byref(self.ctx), self._seq_put[dst], dst, t.data_ptr(),
                                                self.rank * self.MAILBOX_BYTES, nbytes, self._sp()), "comm_put")

    def get(self, t: torch.Tensor, src: int) -> torch.Tensor:
        """K9: wait (on the device) for rank ``src``'s next put, then copy it out of the mailbox."""
        nbytes = t.numel() * t.element_size()
        self._seq_get[src] += 1
        _native.check(self.lib.b200rnn_comm_wait(C.byref(self.ctx), self._seq_get[src], src, self._sp()), "comm_wait")
        off = src * self.MAILBOX_BYTES
        t.view(-1).view(torch.uint8).copy_(self.heap[off:off + nbytes])
        return t

    # ---- K10: one-sided parameter server primitives ----
    def ps_pull(self, dst: torch.Tensor, server: int, heap_off: int) -> None:
        _native.check(self.lib.b200rnn_comm_ps_pull(C.byref(self.ctx), server, heap_off, dst.data_ptr(), dst.numel(),
                                                    self._sp()), "comm_ps_pull")
        _native.count_launch()

    def ps_push(self, rows: torch.Tensor, n_rows: int, n_params: int, opt, optdesc, metrics_slot, server: int,
                heap_off: int) -> None:
        _native.check(self.lib.b200rnn_comm_ps_push(C.byref(self.ctx), server, heap_off, rows.data_ptr(), n_rows, n_params,
                                                    opt.exp_avg.data_ptr(), opt.exp_avg_sq.data_ptr(), C.byref(optdesc),
                                                    _native.ptr(metrics_slot), self._sp()), "comm_ps_push")
        _native.count_launch()

    def signal_value(self, src: int) -> int:
        """Host read of the K9 signal counter written by rank ``src`` (synchronises)."""
        flags = torch.as_tensor(_External(self.local_base + self.ctx.off_flags_sig, 4 * _native.MAX_PEERS),
                                device=self.device).view(torch.int32)
        return int(flags[src].item())

    def check(self) -> None:
        """Raise if a watchdog fired (a peer's flag never arrived)."""
        st = C.c_int32(0)
        _native.check(self.lib.b200rnn_comm_status(C.byref(self.ctx), C.byref(st), self._sp()), "comm_status")
        if st.value != 0:
            missing = [r for r in range(self.world) if st.value & (1 << r)]
            extra = " (local grid barrier timed out)" if st.value & (1 << 30) else ""
            raise RuntimeError(f"rank {self.rank}: comm watchdog fired, no signal from rank(s) {missing}{extra}")

    def set_delay_ns(self, ns: int) -> None:
        self.ctx.delay_ns = int(ns)

    def set_loss(self, percent: float, rto_ns: int = 0) -> None:
        """Injected message loss (the reference's `tc netem loss N%`, fabfile.py:149-151): every block's publication is
        "dropped" with this probability and then pays `rto_ns` (0 = 200 us) per retransmission before it goes out."""
        if not 0.0 <= percent <= 100.0:
            raise ValueError("loss percentage must be within [0, 100]")
        self.ctx.loss_ppm = int(round(percent * 10000))
        self.ctx.loss_rto_ns = int(rto_ns)

    def close(self) -> None:
        if self.symm:
            torch.cuda.synchronize(self.device)
            if self.world > 1:
                self.boot.barrier()
            if self._vmm:
                self.lib.b200rnn_symm_vmm_destroy(self.symm)
            else:
                self.lib.b200rnn_symm_destroy(self.symm)
            self.symm = C.c_void_p()
        self.boot.close()
