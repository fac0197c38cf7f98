"""Request batcher; drop-in for the reference's Batcher (src/batcher.py).

Buffering, the size / latency flush triggers and the future fan-out are host-side asyncio, as in
the reference.  What moves to the GPU is the data path of ``_process_batch``:

  PACK   ``inputs_list = [req.inputs for req in requests]``            (src/batcher.py:211)
         -> K4 ``die_pack_ragged``: ragged payloads -> contiguous padded ``[B, Lpad]`` device batch
  INFER  ``await batch_callback(model, version, inputs_list)``         (src/batcher.py:212-216)
         -> ``device_callback(model, version, PackedBatch) -> PackedBatch``
  UNPACK ``for req, result in zip(...): req.future.set_result(result)``  (src/batcher.py:225-227)
         -> K5 ``die_unpack_ragged``: padded rows -> per-request result slices

With a plain ``batch_callback`` (arbitrary Python objects) the callback contract is exactly the
reference's and no device is involved; ``device_callback`` (keyword-only, additive) selects the
kernel path for numeric payloads.  Row i of the packed batch is always the i-th request in arrival
order for its ``model:version`` key, and result i goes to request i.
"""
from __future__ import annotations

import asyncio
import contextlib
import logging
import time
import uuid
from dataclasses import dataclass, field
from typing import Any, Awaitable, Callable, Dict, List, Optional

logger = logging.getLogger(__name__)


@dataclass
class BatchedRequest:
    request_id: str
    model_name: str
    version: str
    inputs: Any
    created_at: float
    future: asyncio.Future = field(default_factory=lambda: asyncio.Future())


@dataclass
class Batch:
    model_name: str
    version: str
    requests: List[BatchedRequest]
    created_at: float
    max_batch_size: int
    max_latency: float


@dataclass
class PackedBatch:
    """Contiguous padded device batch: ``data`` [B, Lpad] (CUDA), ``lengths`` int32 [B] (CUDA) and the
    same lengths on the host (known at pack time, so unpack needs no device read-back)."""
    data: Any
    lengths: Any
    host_lengths: List[int]


class Batcher:
    def __init__(self, max_batch_size: int = 32, max_latency_ms: float = 100.0,
                 batch_callback: Optional[Callable[[str, str, List[Any]], Awaitable[List[Any]]]] = None,
                 *, device_callback: Optional[Callable[[str, str, PackedBatch], Awaitable[PackedBatch]]] = None):
        if max_batch_size < 1:
            raise ValueError("max_batch_size must be at least 1")
        if max_latency_ms <= 0:
            raise ValueError("max_latency_ms must be greater than 0")
        self.max_batch_size = max_batch_size
        self.max_latency = max_latency_ms / 1000.0
        self.batch_callback = batch_callback
        self.device_callback = device_callback
        self._batches: Dict[str, Batch] = {}
        self._flush_tasks: Dict[str, asyncio.Task] = {}
        self._running = False
        self._lock = asyncio.Lock()
        self.total_batches = 0
        self.total_requests = 0
        self.total_batched_requests = 0

    async def start(self) -> None:
        self._running = True

    async def stop(self) -> None:
        if not self._running:
            return
        self._running = False
        async with self._lock:
            timers = list(self._flush_tasks.values())
            leftovers = list(self._batches.items())
            self._flush_tasks.clear()
            self._batches.clear()
        for t in timers:
            if not t.done():
                t.cancel()
                with contextlib.suppress(asyncio.CancelledError):
                    await t
        for key, batch in leftovers:
            if batch.requests:
                try:
                    await self._process_batch(key, batch, list(batch.requests))
                except Exception:  # noqa: BLE001
                    logger.exception("Error processing batch %s during shutdown", key)

    # -- ingest (src/batcher.py:102-149) ---------------------------------------------------------------
    async def add_request(self, model_name: str, version: str, inputs: Any,
                          request_id: Optional[str] = None) -> asyncio.Future:
        if not self._running:
            raise RuntimeError("Batcher is not running")
        req = BatchedRequest(request_id=request_id or str(uuid.uuid4()), model_name=model_name, version=version,
                             inputs=inputs, created_at=time.time())
        key = f"{model_name}:{version}"
        async with self._lock:
            batch = self._batches.get(key)
            if batch is None:
                batch = Batch(model_name=model_name, version=version, requests=[], created_at=time.time(),
                              max_batch_size=self.max_batch_size, max_latency=self.max_latency)
                self._batches[key] = batch
                self._flush_tasks[key] = asyncio.create_task(self._schedule_flush(key, batch))
            batch.requests.append(req)
            self.total_requests += 1
            full = len(batch.requests) >= self.max_batch_size
        if full:
            await self._flush_batch(key, batch)  # runs inline in the producer, like the reference
        return req.future

    add = add_request  # north-star alias

    async def flush(self, key: Optional[str] = None) -> None:
        """Flush one ``model:version`` buffer (or all) now; additive helper, not in the reference."""
        async with self._lock:
            todo = [(k, b) for k, b in self._batches.items() if key is None or k == key]
        for k, b in todo:
            await self._flush_batch(k, b)

    async def _schedule_flush(self, key: str, batch: Batch) -> None:
        try:
            await asyncio.sleep(batch.max_latency)
            async with self._lock:  # the reference processes the timed-out batch while holding the lock
                if self._batches.get(key) is batch and batch.requests:
                    reqs = list(batch.requests)
                    batch.requests.clear()
                    del self._batches[key]
                    await self._process_batch(key, batch, reqs)
        except asyncio.CancelledError:
            pass
        except Exception:  # noqa: BLE001
            logger.exception("Unexpected error in scheduled flush for batch %s", key)
            async with self._lock:
                self._batches.pop(key, None)
                self._flush_tasks.pop(key, None)

    async def _flush_batch(self, key: str, batch: Batch) -> None:
        async with self._lock:
            if not batch.requests:
                return
            reqs = list(batch.requests)
            batch.requests.clear()
            if self._batches.get(key) is batch:
                del self._batches[key]
            timer = self._flush_tasks.pop(key, None)
            if timer is not None and not timer.done():
                timer.cancel()
        await self._process_batch(key, batch, reqs)

    # -- pack / infer / unpack (src/batcher.py:202-240) ---------------------------------------------------
    async def _process_batch(self, key: str, batch: Batch, reqs: List[BatchedRequest]) -> None:
        if not reqs:
            return
        self.total_batches += 1
        self.total_batched_requests += len(reqs)
        try:
            if self.device_callback is not None:
                results = await self._run_device(batch, reqs)
            elif self.batch_callback:
                inputs_list = [r.inputs for r in reqs]
                results = await self.batch_callback(batch.model_name, batch.version, inputs_list)
            else:
                err = RuntimeError("No batch callback configured")
                for r in reqs:
                    if not r.future.done():
                        r.future.set_exception(err)
                return
            if len(results) != len(reqs):
                raise ValueError(f"Batch callback returned {len(results)} results but expected {len(reqs)}")
            for r, res in zip(reqs, results):
                if not r.future.done():
                    r.future.set_result(res)
        except Exception as e:  # noqa: BLE001 - every future of the batch sees the failure
            logger.exception("Error processing batch %s", key)
            for r in reqs:
                if not r.future.done():
                    r.future.set_exception(e)

    async def _run_device(self, batch: Batch, reqs: List[BatchedRequest]) -> List[Any]:
        packed = pack_requests([r.inputs for r in reqs])
        out = await self.device_callback(batch.model_name, batch.version, packed)
        return unpack_results(out)

    async def get_stats(self) -> Dict[str, Any]:
        avg = self.total_batched_requests / self.total_batches if self.total_batches > 0 else 0
        async with self._lock:
            pending_requests = sum(len(b.requests) for b in self._batches.values() if b.requests)
        return {"total_batches": self.total_batches, "total_requests": self.total_requests,
                "total_batched_requests": self.total_batched_requests, "pending_batches": len(self._batches),
                "pending_requests": pending_requests, "avg_batch_size": avg, "max_batch_size": self.max_batch_size,
                "max_latency_ms": self.max_latency * 1000.0}


def pack_requests(inputs: List[Any], dtype=None, Lpad: Optional[int] = None) -> PackedBatch:
    """PACK on the device: per-request payloads (1-D CUDA tensors, numpy arrays or number lists)
    -> ``PackedBatch``.  Host payloads are staged as one CSR buffer and copied once."""
    import numpy as np
    import torch
    from . import ops
    if all(isinstance(x, torch.Tensor) and x.is_cuda for x in inputs):
        lens = [int(x.numel()) for x in inputs]
        values = torch.cat([x.reshape(-1) for x in inputs]) if sum(lens) else torch.zeros(
            4, dtype=dtype or inputs[0].dtype, device="cuda")
    else:
        want = np.dtype(dtype or np.float32)
        if type(inputs[0]) is np.ndarray and inputs[0].dtype == want and inputs[0].ndim == 1 and \
                type(inputs[-1]) is np.ndarray and inputs[-1].ndim == 1:
            arrs = inputs  # the usual case: numpy rows as they are (np.concatenate casts a stray dtype, rejects a stray shape)
        else:
            arrs = [np.asarray(x, dtype=want).reshape(-1) for x in inputs]
        lens = list(map(len, arrs))
        total = sum(lens)
        # one concatenate straight into a reusable pinned buffer, one H2D copy
        stage = ops.pinned_staging("payload", max(total, 4) * want.itemsize)
        host = stage.numpy()[:max(total, 4) * want.itemsize].view(want)
        if total:
            np.concatenate(arrs, out=host[:total])
        else:
            host[:] = 0
        values = torch.empty(max(total, 4), dtype=torch.from_numpy(host[:1]).dtype, device="cuda")
        values.copy_(torch.from_numpy(host), non_blocking=True)
        torch.cuda.current_stream().synchronize()  # the staging buffer is reused by the next call
    off = np.zeros(len(lens) + 1, dtype=np.int64)
    np.cumsum(lens, out=off[1:])
    per16 = max(16 // values.element_size(), 1)
    width = Lpad if Lpad is not None else max(max(lens, default=0), 1)
    width = (width + per16 - 1) // per16 * per16  # rows are whole 16-byte vectors
    data, lengths = ops.pack_ragged(values, torch.from_numpy(off).to("cuda"), width)
    return PackedBatch(data=data, lengths=lengths, host_lengths=lens)


def unpack_results(packed: PackedBatch) -> List[Any]:
    """UNPACK on the device: padded rows -> one flat result buffer; returns per-request views."""
    import numpy as np
    import torch
    from . import ops
    lens = packed.host_lengths
    off = np.zeros(len(lens) + 1, dtype=np.int64)
    np.cumsum(lens, out=off[1:])
    flat = torch.empty(max(int(off[-1]), 1), dtype=packed.data.dtype, device=packed.data.device)
    ops.unpack_ragged(packed.data, torch.from_numpy(off).to(packed.data.device), flat)
    return [flat[off[i]:off[i + 1]] for i in range(len(lens))]
