"""ctypes binding of ``libb200rnn.so`` (C ABI in ``include/b200rnn.h``).

The library is built in-tree by :mod:`._build` and loaded from
``pytorch_distributed_rnn_b200/_lib``.  On a CUDA machine a missing or unloadable
library is a hard error (no silent fallback to eager PyTorch); on a CPU-only
machine the package works without it through the torch reference path.
"""
from __future__ import annotations

import ctypes as C
import threading
from typing import Optional

import torch

from . import _build

MAX_LAYERS = 8
MAX_PEERS = 16

_lib: Optional[C.CDLL] = None
_lib_err: Optional[str] = None
_lock = threading.Lock()


class ModelDesc(C.Structure):
    _fields_ = [
        ("cell", C.c_int32), ("input_size", C.c_int32), ("hidden_size", C.c_int32),
        ("num_layers", C.c_int32), ("batch", C.c_int32), ("seq_len", C.c_int32),
        ("num_classes", C.c_int32), ("reserved", C.c_int32),
        ("off_wih", C.c_int64 * MAX_LAYERS), ("off_whh", C.c_int64 * MAX_LAYERS),
        ("off_bih", C.c_int64 * MAX_LAYERS), ("off_bhh", C.c_int64 * MAX_LAYERS),
        ("off_fcw", C.c_int64), ("off_fcb", C.c_int64), ("n_params", C.c_int64),
    ]


class OptimDesc(C.Structure):
    _fields_ = [
        ("kind", C.c_int32), ("step", C.c_int32),
        ("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
        ("weight_decay", C.c_float), ("momentum", C.c_float), ("grad_scale", C.c_float),
        ("reserved", C.c_int32),
    ]


class CommCtx(C.Structure):
    _fields_ = [
        ("rank", C.c_int32), ("world", C.c_int32), ("max_blocks", C.c_int32), ("delay_ns", C.c_int32),
        ("inbox_floats", C.c_int64), ("timeout_cycles", C.c_int64), ("heap_bytes", C.c_int64),
        ("ll_floats", C.c_int64), ("loss_ppm", C.c_int32), ("loss_rto_ns", C.c_int32),
        ("off_flags_ar", C.c_int64), ("off_flags_pr", C.c_int64), ("off_flags_sig", C.c_int64),
        ("off_flags_bar", C.c_int64), ("off_status", C.c_int64), ("off_inbox", C.c_int64), ("off_heap", C.c_int64),
        ("off_ll", C.c_int64),
        ("base", C.c_void_p * MAX_PEERS),
    ]


class Topology(C.Structure):
    _fields_ = [
        ("n_devices", C.c_int32),
        ("sm_count", C.c_int32 * MAX_PEERS),
        ("multicast_supported", C.c_int32 * MAX_PEERS),
        ("can_access", (C.c_int32 * MAX_PEERS) * MAX_PEERS),
        ("perf_rank", (C.c_int32 * MAX_PEERS) * MAX_PEERS),
        ("native_atomics", (C.c_int32 * MAX_PEERS) * MAX_PEERS),
        ("name", (C.c_char * 64) * MAX_PEERS),
    ]


class NativeError(RuntimeError):
    pass


def _declare(lib: C.CDLL) -> None:
    vp, i32, i64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float
    P = C.POINTER
    lib.b200rnn_last_error.restype = C.c_char_p
    lib.b200rnn_version.restype = i32
    sigs = {
        "b200rnn_gather_batch": (i32, [vp, vp, vp, i32, i32, vp, vp, vp]),
        "b200rnn_gather_batch_capped": (i32, [vp, vp, vp, i32, i32, vp, vp, i32, vp]),
        "b200rnn_rnn_small_supported": (i32, [P(ModelDesc)]),
        "b200rnn_rnn_small_tiling": (i32, [P(ModelDesc), P(i32), P(i32)]),
        "b200rnn_rnn_stash_floats": (i64, [P(ModelDesc)]),
        "b200rnn_rnn_bwd_scratch_floats": (i64, [P(ModelDesc)]),
        "b200rnn_grad_row_floats": (i64, [P(ModelDesc)]),
        "b200rnn_lstm_fwd": (i32, [P(ModelDesc), vp, vp, vp, i32, vp]),
        "b200rnn_lstm_fwd_tc": (i32, [P(ModelDesc), vp, vp, vp, i32, vp]),
        "b200rnn_tc_status": (i32, [P(i32)]),
        "b200rnn_lstm_bwd": (i32, [P(ModelDesc), vp, vp, vp, vp, f32, vp, vp, vp, vp, vp]),
        "b200rnn_gru_fwd": (i32, [P(ModelDesc), vp, vp, vp, i32, vp]),
        "b200rnn_gru_bwd": (i32, [P(ModelDesc), vp, vp, vp, vp, f32, vp, vp, vp, vp]),
        "b200rnn_lstm_fwd_ex": (i32, [P(ModelDesc), vp, vp, vp, vp, i32, i32, vp]),
        "b200rnn_lstm_bwd_ex": (i32, [P(ModelDesc), vp, vp, vp, vp, vp, f32, vp, vp, vp, vp, i32, vp]),
        "b200rnn_gru_fwd_ex": (i32, [P(ModelDesc), vp, vp, vp, vp, i32, i32, vp]),
        "b200rnn_gru_bwd_ex": (i32, [P(ModelDesc), vp, vp, vp, vp, vp, f32, vp, vp, vp, i32, vp]),
        "b200rnn_rnn_persist_supported": (i32, [P(ModelDesc)]),
        "b200rnn_persist_status": (i32, [P(i32)]),
        "b200rnn_wave_status": (i32, [P(i32)]),
        "b200rnn_rnn_small_fwd_wave": (i32, [P(ModelDesc)]),
        "b200rnn_rnn_small_bwd_wave": (i32, [P(ModelDesc)]),
        "b200rnn_lstm_bwd_dx": (i32, [P(ModelDesc), vp, vp, vp, vp, f32, vp, vp, vp, vp, vp, vp]),
        "b200rnn_gru_bwd_dx": (i32, [P(ModelDesc), vp, vp, vp, vp, f32, vp, vp, vp, vp, vp]),
        "b200rnn_lstm_bwd_static": (i32, [P(ModelDesc), vp, vp, vp, vp, f32, vp, vp, vp, vp, vp]),
        "b200rnn_reduce_rows": (i32, [vp, i32, i64, vp, i32, vp]),
        "b200rnn_head_fwd_metrics_floats": (i64, [i32]),
        "b200rnn_head_fwd": (i32, [vp, i64, vp, vp, vp, i32, i32, i32, vp, vp, vp]),
        "b200rnn_fused_update": (i32, [vp, i32, i64, vp, vp, vp, P(OptimDesc), vp, vp, vp]),
        "b200rnn_l2_flush": (i32, [vp, i64, vp]),
        "b200rnn_rnn_large_stash_bytes": (i64, [P(ModelDesc)]),
        "b200rnn_rnn_large_scratch_bytes": (i64, [P(ModelDesc)]),
        "b200rnn_rnn_large_h_offset": (i64, [P(ModelDesc), i32]),
        "b200rnn_rnn_large_c_offset": (i64, [P(ModelDesc), i32]),
        "b200rnn_rnn_large_fwd": (i32, [P(ModelDesc), vp, vp, vp, vp, vp]),
        "b200rnn_rnn_large_bwd": (i32, [P(ModelDesc), vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]),
        "b200rnn_gemm_bf16": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp]),
        "b200rnn_gemm_status": (i32, [P(i32)]),
        "b200rnn_cast_bf16": (i32, [vp, vp, i64, vp]),
        "b200rnn_transpose_cast_bf16": (i32, [vp, vp, i32, i32, i64, i64, vp]),
        # runtime bootstrap / symmetric memory / reducer
        "b200rnn_boot_listen": (i32, [P(i32), P(i32)]),
        "b200rnn_boot_init": (i32, [i32, i32, C.c_char_p, i32, i32, i32, P(vp)]),
        "b200rnn_boot_allgather": (i32, [vp, vp, i32, vp]),
        "b200rnn_boot_barrier": (i32, [vp]),
        "b200rnn_boot_destroy": (None, [vp]),
        "b200rnn_topology_probe": (i32, [P(Topology)]),
        "b200rnn_symm_create": (i32, [vp, i32, i64, P(vp)]),
        "b200rnn_symm_pointers": (i32, [vp, P(vp)]),
        "b200rnn_symm_destroy": (i32, [vp]),
        "b200rnn_symm_vmm_create": (i32, [vp, i32, i32, i32, i64, i32, i32, P(vp)]),
        "b200rnn_symm_vmm_pointers": (i32, [vp, P(vp), P(vp)]),
        "b200rnn_symm_vmm_destroy": (i32, [vp]),
        "b200rnn_bucket_plan": (i32, [P(i64), i32, i64, i64, P(i32), P(i32)]),
        "b200rnn_reducer_create": (i32, [i32, P(i64), P(i64), P(vp)]),
        "b200rnn_reducer_destroy": (i32, [vp]),
        "b200rnn_reducer_range": (i32, [vp, i32, P(i64), P(i64)]),
        "b200rnn_reducer_counter": (vp, [vp, i32]),
        "b200rnn_reducer_expect": (i32, [vp, i32, i32]),
        "b200rnn_reducer_wait": (i32, [vp, i32, vp]),
        "b200rnn_reducer_stamp": (i32, [vp, i32, i32, vp]),
        "b200rnn_reducer_read": (i32, [vp, P(C.c_uint64), P(i32)]),
        "b200rnn_head_ce_scratch_floats": (i64, [i32, i32]),
        "b200rnn_head_ce": (i32, [vp, i64, vp, vp, vp, vp, i32, i32, i32, f32, vp, i64, vp, vp, vp, vp, vp, P(i32), vp]),
        "b200rnn_head_status": (i32, [P(i32)]),
        "b200rnn_gather_batch_tm": (i32, [vp, vp, i32, i32, i32, vp, vp]),
        "b200rnn_gemm_bf16_retire": (i32, [vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp, P(i32), vp]),
        "b200rnn_rnn_large_bwd_layer": (i32, [P(ModelDesc), vp, vp, vp, vp, i32, vp, vp, vp, vp, vp, vp, P(i32), vp]),
        # in-kernel collectives
        "b200rnn_comm_layout": (i64, [P(CommCtx)]),
        "b200rnn_comm_allreduce_update": (i32, [P(CommCtx), C.c_uint32, vp, i32, i64, vp, i64, vp, vp, P(OptimDesc), vp, i32, vp]),
        "b200rnn_comm_allreduce_update_nvls": (i32, [P(CommCtx), vp, C.c_uint32, vp, i32, i64, vp, i64, i64, vp, vp, P(OptimDesc), vp, vp]),
        "b200rnn_comm_allreduce": (i32, [P(CommCtx), C.c_uint32, vp, i64, f32, vp]),
        "b200rnn_comm_allreduce_algo": (i32, [P(CommCtx), C.c_uint32, vp, i64, f32, i32, vp]),
        "b200rnn_comm_broadcast": (i32, [P(CommCtx), C.c_uint32, i64, i64, i32, vp]),
        "b200rnn_comm_put": (i32, [P(CommCtx), C.c_uint32, i32, vp, i64, i64, vp]),
        "b200rnn_comm_wait": (i32, [P(CommCtx), C.c_uint32, i32, vp]),
        "b200rnn_comm_ps_push": (i32, [P(CommCtx), i32, i64, vp, i32, i64, vp, vp, P(OptimDesc), vp, vp]),
        "b200rnn_comm_ps_pull": (i32, [P(CommCtx), i32, i64, vp, i64, vp]),
        "b200rnn_comm_forget": (i32, [vp]),
        "b200rnn_comm_barrier": (i32, [P(CommCtx), C.c_uint32, vp]),
        "b200rnn_comm_status": (i32, [P(CommCtx), P(i32), vp]),
    }
    for name, (res, args) in sigs.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args


def load(required: bool = False) -> Optional[C.CDLL]:
    """Load (building first if stale and nvcc is present) the native library."""
    global _lib, _lib_err
    with _lock:
        if _lib is not None:
            return _lib
        if _lib_err is None:
            try:
                if _build.is_stale() and _build.find_nvcc() is not None:
                    _build.build()
                lib = C.CDLL(str(_build.LIB_PATH))
                _declare(lib)
                _lib = lib
                return _lib
            except (OSError, RuntimeError, AttributeError) as e:  # missing nvcc / .so / symbol
                _lib_err = f"{type(e).__name__}: {e}"
        if required:
            raise NativeError(
                "libb200rnn.so is not available (%s). Build it with "
                "`python -m pytorch_distributed_rnn_b200._build`." % _lib_err)
        return None


def available() -> bool:
    return load() is not None


def lib() -> C.CDLL:
    return load(required=True)


def device_status() -> dict:
    """Watchdog words of the kernels with bounded in-kernel waits (synchronises the device): the tcgen05 GEMM
    pipeline, the persistent large-hidden time loops and the small-hidden layer wavefront.  All zero = healthy."""
    out = {}
    l = lib()
    for name in ("gemm", "persist", "wave", "tc", "head"):
        fn = getattr(l, f"b200rnn_{name}_status", None)
        if fn is None:
            continue
        st = C.c_int32(0)
        check(fn(C.byref(st)), f"{name}_status")
        out[name] = st.value
    return out


def assert_device_healthy() -> None:
    """Raise if any in-kernel watchdog fired since the library was loaded (results of that call are garbage)."""
    bad = {k: v for k, v in device_status().items() if v != 0}
    if bad.get("wave", 0) & 2 or bad.get("head", 0) & 1:
        raise NativeError("fused head: a label outside [0, num_classes) reached the loss (status %r)" % (bad,))
    if bad:
        raise NativeError(f"in-kernel watchdog fired: {bad} (a wait for a peer CTA / TMA / MMA timed out; "
                          "see INTEGRATION.md: B200RNN_PERSIST / B200RNN_WAVE)")


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().b200rnn_last_error().decode("utf-8", "replace")
        raise NativeError(f"{what or 'b200rnn call'} failed (rc={rc}): {msg}")


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    """Raw device pointer of a tensor (None -> NULL)."""
    return None if t is None else t.data_ptr()


def stream_ptr(device: Optional[torch.device] = None) -> int:
    return torch.cuda.current_stream(device).cuda_stream


# ---- launch accounting (bench.py reports it as "gpu_launches") ----
_launches = 0


def count_launch(n: int = 1) -> None:
    global _launches
    _launches += n


def launches() -> int:
    return _launches
