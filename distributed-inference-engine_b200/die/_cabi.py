"""ctypes binding of libdie.so (declared in include/die.h).

The product path has no CPU fallback: if the shared library is missing, or a compute entry
point is called without a CUDA device, this module raises.  Loading the library itself does
not need a GPU (the CPU-only test tier checks the exported symbols this way).
"""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_int32, c_int64, c_uint8, c_uint32, c_uint64, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("DIE_LIB_PATH") or os.path.join(_HERE, "_lib", "libdie.so")  # override: A/B diagnostics only

DIE_OK, DIE_ERR_INVALID, DIE_ERR_CUDA, DIE_ERR_NOMEM, DIE_ERR_STOP_ITERATION, DIE_ERR_TOO_LARGE, \
    DIE_ERR_UNSUPPORTED, DIE_ERR_VALUE_ERROR = range(8)
POLICY_IDS = {"lru": 0, "lfu": 1, "fifo": 2}
LINEAR_FP32, LINEAR_TF32 = 0, 1
MODEL_ECHO, MODEL_LINEAR_FP32, MODEL_LINEAR_TF32, MODEL_MLP_BF16 = 0, 1, 2, 3


class DieError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"libdie status {status}: {message}")
        self.status = status


class FlushDesc(ctypes.Structure):  # die_flush_desc_t
    _fields_ = [("queue", c_int32), ("count", c_int32), ("trigger", c_int32), ("pad", c_int32), ("opened_at", c_double)]


class PipelineConfig(ctypes.Structure):
    _fields_ = [("kv_max_size", c_uint64), ("default_ttl", c_double), ("max_batch", c_int32),
                ("dim_in", c_int32), ("dim_out", c_int32), ("model", c_int32), ("max_key_bytes", c_int32),
                ("hidden", c_int32), ("weight", c_void_p), ("bias", c_void_p)]


P = c_void_p  # device or host pointer passed as an integer address

# name -> (restype, argtypes); mirrors include/die.h one to one
SIGNATURES = {
    "die_abi_version": (c_int32, []),
    "die_last_error": (c_char_p, []),
    "die_set_device": (c_int32, [c_int32]),
    "die_launch_count": (c_uint64, []),
    "die_replay_count": (c_uint64, []),
    "die_profile_begin": (c_int32, []),
    "die_profile_end": (c_int32, [c_int32, POINTER(c_double), POINTER(c_uint64), POINTER(c_int32)]),
    "die_debug_trace_enable": (c_int32, [c_uint64]),
    "die_debug_trace_read_sync": (c_int32, [c_uint64, POINTER(c_uint64), POINTER(c_uint64), POINTER(c_uint64)]),
    "die_profile_timeline": (c_int32, [c_int32, POINTER(c_int32), POINTER(c_int32), POINTER(c_double), POINTER(c_double),
                                       POINTER(c_int32)]),
    "die_profile_stage_name": (c_char_p, [c_int32]),
    "die_debug_random_read": (c_int32, [P, c_uint64, c_int32, c_int32, c_int32, c_uint64, P, P]),
    "die_route_md5": (c_int32, [P, P, c_int32, c_int32, P, P, c_int32, c_int32, P, P, P, P]),
    "die_bucket_by_shard": (c_int32, [P, c_int32, c_int32, P, P, P, P]),
    "die_bucket_scatter_rows": (c_int32, [P, P, P, P, c_int32, c_int32, c_int32, c_int32, P, P, P]),
    "die_bucket_gather_rows": (c_int32, [P, P, P, P, c_int32, c_int32, c_int32, c_int32, P, P, P, P]),
    "die_kv_create": (c_int32, [c_uint64, c_int32, c_double, POINTER(c_void_p)]),
    "die_kv_destroy": (c_int32, [P]),
    "die_kv_get_batch": (c_int32, [P, P, c_int32, c_double, P, P, P, P]),
    "die_kv_set_batch": (c_int32, [P, P, P, P, c_int32, c_double, c_int32, P, P, P]),
    "die_kv_set_batch_noalloc": (c_int32, [P, P, P, P, c_int32, c_double, c_int32, P, P, P]),
    "die_kv_delete_batch": (c_int32, [P, P, c_int32, P, P, P]),
    "die_kv_contains_batch": (c_int32, [P, P, c_int32, c_double, P, P, P]),
    "die_kv_len_sync": (c_int32, [P, c_double, P, c_uint64, POINTER(c_uint64), POINTER(c_uint64), P]),
    "die_kv_clear": (c_int32, [P, P]),
    "die_kv_stats_sync": (c_int32, [P, POINTER(c_uint64), P]),
    "die_kv_sync_status": (c_int32, [P, POINTER(c_int32), POINTER(c_uint64), POINTER(c_uint32), P]),
    "die_kv_dump_sync": (c_int32, [P, c_uint64, P, P, P, P, P, P, P, POINTER(c_uint64), P, P]),
    "die_kv_snapshot_bytes": (c_uint64, [P]),
    "die_kv_snapshot_sync": (c_int32, [P, P, c_uint64, P]),
    "die_kv_restore_sync": (c_int32, [P, P, c_uint64, P]),
    "die_kv_maintain_sync": (c_int32, [P, c_int32, P]),
    "die_pack_ragged": (c_int32, [P, P, P, c_int32, c_int32, c_int32, P, P, P]),
    "die_unpack_ragged": (c_int32, [P, P, P, c_int32, c_int32, c_int32, P, P]),
    "die_echo_fwd": (c_int32, [P, P, c_uint64, P]),
    "die_linear_fwd_f32": (c_int32, [P, P, P, P, c_int32, c_int32, c_int32, c_int32, P]),
    "die_mlp_fwd_bf16": (c_int32, [P, P, P, P, P, P, P, c_int32, c_int32, c_int32, P]),
    "die_pipeline_create": (c_int32, [POINTER(PipelineConfig), POINTER(c_void_p)]),
    "die_pipeline_destroy": (c_int32, [P]),
    "die_pipeline_kv": (c_void_p, [P]),
    "die_pipeline_debug_marks_sync": (c_int32, [P, POINTER(c_uint64)]),
    "die_pipeline_debug_cta_marks_sync": (c_int32, [P, c_int32, POINTER(c_uint64)]),
    "die_pipeline_run_device": (c_int32, [P, P, P, c_int32, P, c_double, P, P, P, P]),
    "die_pipeline_run_owner_device": (c_int32, [P, P, c_int32, P, c_double, P, P, P]),
    "die_pipeline_run_owner_device_n": (c_int32, [P, P, c_int32, P, P, c_double, P, P, P]),
    "die_xchg_dispatch": (c_int32, [P, P, P, P, c_int32, c_int32, c_int32, c_int32, c_int32, P, P, P]),
    "die_xchg_compact": (c_int32, [P, P, c_int32, c_int32, c_int32, P, P, P, P, P]),
    "die_xchg_return": (c_int32, [P, P, P, c_int32, c_int32, c_int32, c_int32, P, P]),
    "die_xchg_scatter": (c_int32, [P, P, P, c_int32, c_int32, c_int32, c_int32, P, P, P]),
    "die_xchg_window_bytes": (c_uint64, [c_int32, c_int32, c_int32]),
    "die_xchg_create": (c_int32, [c_int32, c_int32, c_int32, c_int32, c_int32, P, P, POINTER(c_void_p)]),
    "die_xchg_destroy": (c_int32, [P]),
    "die_xchg_step": (c_int32, [P, P, P, c_int32, P, c_double, P, P, P, P]),
    "die_xchg_step_many": (c_int32, [P, c_int32, P, c_int64, P, c_int64, c_int32, P, c_double, c_double, P, P, c_int32, P]),
    "die_xchg_check_sync": (c_int32, [P, P]),
    "die_xchg_prepare": (c_int32, [P, c_int32]),
    "die_xchg_set_health": (c_int32, [P, P, P, c_int32, c_int32, P]),
    "die_frame_decode": (c_int32, [P, P, c_int32, c_int32, P, P, P, P, P, P, P, P]),
    "die_batcher_create": (c_int32, [c_int32, c_int32, c_double, c_int32, c_int32, c_int32, P, P, POINTER(c_void_p)]),
    "die_batcher_destroy": (c_int32, [P]),
    "die_batcher_tick": (c_int32, [P, c_double, c_int32, P, P, P, c_int32, P, P, P]),
    "die_batcher_flushes_sync": (c_int32, [P, POINTER(c_int32), P, P, c_int32, POINTER(c_uint64), P]),
    "die_batcher_stats_sync": (c_int32, [P, POINTER(c_uint64), POINTER(c_uint64), POINTER(c_uint64), P, P]),
    "die_pipeline_run_host": (c_int32, [P, P, P, c_int32, P, c_double, P, P, P]),
    "die_pipeline_run_device_many": (c_int32, [P, c_int32, P, c_int64, P, c_int64, c_int32, P, c_double, c_double, P, P,
                                               c_int32, P]),
    "die_pipeline_run_host_many": (c_int32, [P, c_int32, P, c_int64, P, c_int64, c_int32, P, c_double, c_double, P, P, P]),
}

_lib = None


def load() -> ctypes.CDLL:
    """Load libdie.so and bind every symbol of include/die.h; raises if the library is absent."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU fallback for the dispatch path)")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export a declared symbol
        fn.restype, fn.argtypes = res, args
    if lib.die_abi_version() != 1:
        raise ImportError(f"libdie ABI version {lib.die_abi_version()} != 1")
    _lib = lib
    return lib


def check(status: int) -> None:
    if status != DIE_OK:
        raise DieError(status, (load().die_last_error() or b"").decode(errors="replace"))


def require_cuda():
    """Compute entry points need a CUDA device; fail loudly instead of degrading."""
    import torch
    if not torch.cuda.is_available():
        raise RuntimeError("die: no CUDA device visible; the dispatch kernels have no CPU fallback")
    lib = load()
    check(lib.die_set_device(torch.cuda.current_device()))
    return lib


def stream_ptr():
    import torch
    return torch.cuda.current_stream().cuda_stream


def ptr(t):
    """Raw address of a torch tensor (device or pinned host) or None."""
    return None if t is None else t.data_ptr()
