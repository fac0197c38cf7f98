"""Device-resident Batcher: the reference's per-key accumulation with size- and latency-triggered flush
(src/batcher.py:102-200) for requests that already live on the GPU; binds die_batcher_* (csrc/batcher.cu).

The host-side `die.Batcher` mirrors the reference's asyncio class one to one.  This class is the batched form for a
serving loop that holds device tensors: time advances in ticks, `tick(now, ...)` fires the timers that are due, appends
the tick's arrivals (flushing every batch they fill) and, with `force=True`, does what `Batcher.stop()` does.  Flushed
batches run through the model on the device and every result row lands in `out[request_id]`.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _cabi

TRIGGERS = ("latency", "size", "forced")


class DeviceBatcher:
    def __init__(self, keys: Sequence[Tuple[str, str]], max_batch_size: int = 32, max_latency_ms: float = 50.0,
                 dim_in: int = 128, dim_out: Optional[int] = None, max_arrivals: int = 4096,
                 weights: Optional[np.ndarray] = None, bias: Optional[np.ndarray] = None):
        """keys: the (model_name, version) pairs this batcher serves; queue q is keys[q] ("model:version",
        batcher.py:120).  weights [len(keys), dim_out, dim_in] / bias [len(keys), dim_out]: the mock linear model of
        every key; None: FakeModel's echo.  Same argument checks as the reference's constructor (batcher.py:44-47)."""
        if max_batch_size <= 0:
            raise ValueError("max_batch_size must be positive")
        if max_latency_ms <= 0:
            raise ValueError("max_latency_ms must be positive")
        self.lib = _cabi.load()
        self.keys = [f"{m}:{v}" for m, v in keys]
        self.queue_of: Dict[str, int] = {k: i for i, k in enumerate(self.keys)}
        self.max_batch_size, self.max_latency = int(max_batch_size), max_latency_ms / 1000.0
        self.dim_in, self.dim_out = int(dim_in), int(dim_out if dim_out is not None else dim_in)
        w = None if weights is None else np.ascontiguousarray(weights, dtype=np.float32)
        b = None if bias is None else np.ascontiguousarray(bias, dtype=np.float32)
        if w is not None and w.shape != (len(self.keys), self.dim_out, self.dim_in):
            raise ValueError("weights must be [len(keys), dim_out, dim_in]")
        h = ctypes.c_void_p()
        _cabi.check(self.lib.die_batcher_create(len(self.keys), self.max_batch_size, self.max_latency, self.dim_in,
                                                self.dim_out, int(max_arrivals),
                                                None if w is None else w.ctypes.data, None if b is None else b.ctypes.data,
                                                ctypes.byref(h)))
        self._h = h.value

    def __del__(self):
        if getattr(self, "_h", None):
            self.lib.die_batcher_destroy(self._h)
            self._h = None

    def tick(self, now: float, queue: Optional[torch.Tensor], request_id: Optional[torch.Tensor],
             payload: Optional[torch.Tensor], out: torch.Tensor, batch_of: Optional[torch.Tensor] = None,
             force: bool = False) -> None:
        """queue [n] int32, request_id [n] int32, payload [n, dim_in] fp32 (device); out [*, dim_out] fp32 indexed by
        request id; batch_of [*] int32 (optional).  Enqueued on the current stream; nothing is read back."""
        n = 0 if queue is None else int(queue.shape[0])
        ptr = _cabi.ptr
        _cabi.check(self.lib.die_batcher_tick(self._h, float(now), 1 if force else 0, ptr(queue), ptr(request_id), ptr(payload),
                                              n, ptr(out), ptr(batch_of), _cabi.stream_ptr()))

    def last_flushes(self, with_ids: bool = True) -> List[dict]:
        """What the last tick flushed, in flush order: [{"key", "count", "trigger", "opened_at", "batch", "ids"}]."""
        cap = 4096
        n = ctypes.c_int32()
        first = ctypes.c_uint64()
        desc = (_cabi.FlushDesc * cap)()
        ids = np.empty((cap, self.max_batch_size), dtype=np.int32) if with_ids else None
        _cabi.check(self.lib.die_batcher_flushes_sync(self._h, ctypes.byref(n), ctypes.addressof(desc),
                                                      None if ids is None else ids.ctypes.data, cap, ctypes.byref(first),
                                                      _cabi.stream_ptr()))
        res = []
        for j in range(min(n.value, cap)):
            d = desc[j]
            res.append({"key": self.keys[d.queue], "count": d.count, "trigger": TRIGGERS[d.trigger], "opened_at": d.opened_at,
                        "batch": first.value + j, "ids": ids[j, :d.count].tolist() if with_ids else None})
        return res

    def get_stats(self) -> dict:
        """Batcher.get_stats() (batcher.py:242-266), plus "pending": open batch sizes per key."""
        tr, tb, tbr = ctypes.c_uint64(), ctypes.c_uint64(), ctypes.c_uint64()
        pend = (ctypes.c_int32 * len(self.keys))()
        _cabi.check(self.lib.die_batcher_stats_sync(self._h, ctypes.byref(tr), ctypes.byref(tb), ctypes.byref(tbr),
                                                    ctypes.addressof(pend), _cabi.stream_ptr()))
        pending = {k: int(pend[i]) for i, k in enumerate(self.keys) if pend[i] > 0}
        return {"total_batches": tb.value, "total_requests": tr.value, "total_batched_requests": tbr.value,
                "pending_batches": len(pending), "pending_requests": sum(pending.values()),
                "avg_batch_size": (tbr.value / tb.value) if tb.value else 0, "max_batch_size": self.max_batch_size,
                "max_latency_ms": self.max_latency * 1000.0, "pending": pending}
