"""Composed dispatch path for one owner shard: route -> probe -> pack -> infer -> unpack -> fill.

This is the coordinator the reference's README.md:56-60 describes but never wrote (SURVEY.md
call-stack (E)), for fixed-width fp32 payloads.  One call = one flushed batch; every stage is a
kernel of libdie (csrc/pipeline.cu lists them), nothing is computed on the host.
"""
from __future__ import annotations

import ctypes
import math
from typing import Optional, Sequence

import numpy as np
import torch

from . import _cabi, ops
from ._cabi import PipelineConfig, check, ptr, require_cuda, stream_ptr
from .kvstore import DeviceKV

MODEL_IDS = {"echo": _cabi.MODEL_ECHO, "linear": _cabi.MODEL_LINEAR_FP32, "linear_tf32": _cabi.MODEL_LINEAR_TF32,
             "mlp_bf16": _cabi.MODEL_MLP_BF16}


class DispatchPipeline:
    def __init__(self, kv_max_size: int, max_batch: int, dim_in: int, dim_out: Optional[int] = None,
                 model: str = "echo", weight: Optional[np.ndarray] = None, bias: Optional[np.ndarray] = None,
                 default_ttl: Optional[float] = None, max_key_bytes: int = 0):
        self._lib = require_cuda()
        dim_out = dim_in if dim_out is None else dim_out
        self.max_batch, self.dim_in, self.dim_out, self.model = max_batch, dim_in, dim_out, model
        hidden = 0
        if model == "mlp_bf16":  # weight = (W1 [H, D], W2 [D, H]), bias = (b1 [H], b2 [D]): BASELINE.json configs[4]
            w1, w2 = (np.ascontiguousarray(a, dtype=np.float32) for a in weight)
            b1, b2 = (np.ascontiguousarray(a, dtype=np.float32) for a in bias)
            hidden = int(w1.shape[0])
            if w1.shape != (hidden, dim_in) or w2.shape != (dim_out, hidden) or dim_in != dim_out:
                raise ValueError("mlp_bf16 needs W1 [H, D], W2 [D, H] and dim_in == dim_out == D")
            weight, bias = np.concatenate([w1.ravel(), w2.ravel()]), np.concatenate([b1.ravel(), b2.ravel()])
        self._w = None if weight is None else np.ascontiguousarray(weight, dtype=np.float32)
        self._b = None if bias is None else np.ascontiguousarray(bias, dtype=np.float32)
        cfg = PipelineConfig(kv_max_size=kv_max_size, default_ttl=math.nan if default_ttl is None else default_ttl,
                             max_batch=max_batch, dim_in=dim_in, dim_out=dim_out, model=MODEL_IDS[model],
                             max_key_bytes=max_key_bytes, hidden=hidden,
                             weight=None if self._w is None else self._w.ctypes.data,
                             bias=None if self._b is None else self._b.ctypes.data)
        h = ctypes.c_void_p()
        check(self._lib.die_pipeline_create(ctypes.byref(cfg), ctypes.byref(h)))
        self._h = h.value
        self.kv = DeviceKV(kv_max_size, default_ttl, _handle=self._lib.die_pipeline_kv(self._h))

    def __del__(self):
        if getattr(self, "_h", None):
            self._lib.die_pipeline_destroy(self._h)
            self._h = None

    @staticmethod
    def _graph_steps() -> int:
        """Batches per replayed group (environment DIE_GRAPH_STEPS, default 16; 0 = off)."""
        import os
        return int(os.environ.get("DIE_GRAPH_STEPS", "16"))

    def run_device(self, keys: ops.KeyBatch, payload: torch.Tensor, now: float, out: Optional[torch.Tensor] = None,
                   hit: Optional[torch.Tensor] = None):
        """Inputs resident in HBM -> (out [B, dim_out] fp32, hit uint8 [B]); asynchronous on the current stream."""
        B = len(keys)
        out = torch.empty((B, self.dim_out), dtype=torch.float32, device=payload.device) if out is None else out
        hit = torch.empty(B, dtype=torch.uint8, device=payload.device) if hit is None else hit
        check(self._lib.die_pipeline_run_device(self._h, ptr(keys.key_bytes), ptr(keys.key_off), B, ptr(payload), now,
                                                ptr(out), ptr(hit), None, stream_ptr()))
        return out, hit

    def run_owner_device(self, digests: torch.Tensor, payload: torch.Tensor, now: float):
        B = digests.shape[0]
        out = torch.empty((B, self.dim_out), dtype=torch.float32, device=payload.device)
        hit = torch.empty(B, dtype=torch.uint8, device=payload.device)
        check(self._lib.die_pipeline_run_owner_device(self._h, ptr(digests), B, ptr(payload), now, ptr(out), ptr(hit),
                                                      stream_ptr()))
        return out, hit

    def run_host(self, key_bytes: torch.Tensor, key_off: torch.Tensor, payload: torch.Tensor, now: float,
                 out: torch.Tensor, hit: torch.Tensor) -> None:
        """Host (pinned) buffers in, host buffers out; H2D + kernels + D2H + sync inside the call."""
        B = key_off.numel() - 1
        check(self._lib.die_pipeline_run_host(self._h, ptr(key_bytes), ptr(key_off), B, ptr(payload), now, ptr(out),
                                              ptr(hit), stream_ptr()))

    def run_device_many(self, key_bytes: torch.Tensor, key_off: torch.Tensor, payload: torch.Tensor, now0: float,
                        now_step: float, out: torch.Tensor, hit: torch.Tensor) -> None:
        """A queue of flushed batches resident in HBM: key_bytes [n, stride] uint8, key_off [B+1] (shared) or
        [n, B+1], payload [n, B, dim_in]; out [ring, B, dim_out], hit [ring, B] are reused round-robin."""
        n, B = payload.shape[0], payload.shape[1]
        shared = key_off.dim() == 1
        check(self._lib.die_pipeline_run_device_many(self._h, n, ptr(key_bytes), key_bytes.stride(0), ptr(key_off),
                                                     0 if shared else key_off.stride(0), B, ptr(payload), now0, now_step,
                                                     ptr(out), ptr(hit), out.shape[0], stream_ptr()))

    def run_host_many(self, key_bytes: torch.Tensor, key_off: torch.Tensor, payload: torch.Tensor, now0: float,
                      now_step: float, out: torch.Tensor, hit: torch.Tensor) -> None:
        """The same queue through pinned host buffers; out [n, B, dim_out], hit [n, B] (pinned).  Synchronous."""
        n, B = payload.shape[0], payload.shape[1]
        shared = key_off.dim() == 1
        check(self._lib.die_pipeline_run_host_many(self._h, n, ptr(key_bytes), key_bytes.stride(0), ptr(key_off),
                                                   0 if shared else key_off.stride(0), B, ptr(payload), now0, now_step,
                                                   ptr(out), ptr(hit), stream_ptr()))

    def process(self, keys: Sequence[str], payloads: np.ndarray, now: float):
        """Convenience for tests: python strings + numpy rows -> numpy results."""
        kb, off = ops.KeyBatch.encode_host(keys)
        B = len(keys)
        kb_t = torch.from_numpy(np.concatenate([kb, np.zeros(16, np.uint8)])).pin_memory()
        off_t = torch.from_numpy(off).pin_memory()
        pay_t = torch.from_numpy(np.ascontiguousarray(payloads, dtype=np.float32)).pin_memory()
        out = torch.empty((B, self.dim_out), dtype=torch.float32).pin_memory()
        hit = torch.empty(B, dtype=torch.uint8).pin_memory()
        self.run_host(kb_t, off_t, pay_t, now, out, hit)
        return out.numpy().copy(), hit.numpy().copy()
