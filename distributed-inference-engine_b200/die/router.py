"""Hash-shard router with a health gate and failover; drop-in for the reference's Router (src/router.py).

The per-request decision of ``route_request`` (src/router.py:140-184) and ``_find_alternative_shard``
(src/router.py:186-221) is evaluated for a whole batch by the K1 kernel (csrc/route.cu): the host
only lowers the worker table into two tiny device arrays (``routable`` per shard position and the
ordered list of HEALTHY shard positions).  Worker bookkeeping and the TCP health probe are control
plane and stay in Python.
"""
from __future__ import annotations

import asyncio
import logging
import time
from dataclasses import dataclass, field
from enum import Enum, auto
from typing import Any, Dict, List, Optional, Sequence

from .model_registry import ModelRegistry, ModelShard

logger = logging.getLogger(__name__)


class WorkerHealth(Enum):
    HEALTHY = auto()
    UNHEALTHY = auto()
    UNKNOWN = auto()


@dataclass
class WorkerInfo:
    worker_id: str
    address: str
    health: WorkerHealth = WorkerHealth.UNKNOWN
    last_health_check: float = 0.0
    consecutive_failures: int = 0
    last_success: float = 0.0
    metadata: Dict[str, Any] = field(default_factory=dict)


class Router:
    def __init__(self, registry: ModelRegistry, health_check_interval: float = 5.0,
                 health_check_timeout: float = 2.0, max_consecutive_failures: int = 3,
                 failover_enabled: bool = True):
        self.registry = registry
        self.health_check_interval = health_check_interval
        self.health_check_timeout = health_check_timeout
        self.max_consecutive_failures = max_consecutive_failures
        self.failover_enabled = failover_enabled
        self.workers: Dict[str, WorkerInfo] = {}
        self._health_check_task: Optional[asyncio.Task] = None
        self._running = False

    # -- worker table (src/router.py:109-138, 223-245) -----------------------------------------
    def register_worker(self, worker_id: str, address: str, metadata: Optional[Dict[str, Any]] = None) -> None:
        self.workers[worker_id] = WorkerInfo(worker_id=worker_id, address=address, metadata=metadata or {})

    def unregister_worker(self, worker_id: str) -> None:
        self.workers.pop(worker_id, None)

    def mark_worker_success(self, worker_id: str) -> None:
        w = self.workers.get(worker_id)
        if w is not None:
            w.last_success = time.time()
            w.consecutive_failures = 0
            w.health = WorkerHealth.HEALTHY

    def mark_worker_failure(self, worker_id: str) -> None:
        w = self.workers.get(worker_id)
        if w is not None:
            w.consecutive_failures += 1
            if w.consecutive_failures >= self.max_consecutive_failures:
                w.health = WorkerHealth.UNHEALTHY

    # -- the hot op ------------------------------------------------------------------------------
    def health_masks(self, shards: Sequence[ModelShard], prefer_healthy: bool):
        """Lower the worker table to what the kernel consumes: (routable[n] bytes, healthy positions)."""
        routable, healthy = [], []
        for pos, s in enumerate(shards):
            w = self.workers.get(s.worker_id)
            is_healthy = w is not None and w.health == WorkerHealth.HEALTHY
            routable.append(1 if (w is not None and (is_healthy or not prefer_healthy)) else 0)
            if is_healthy:
                healthy.append(pos)
        return routable, healthy

    def route(self, model_name: str, version: str, keys, prefer_healthy: bool = True):
        """Batched route_request.  ``keys``: sequence of str or an ops.KeyBatch already on the device.
        Returns an int32 CUDA tensor of shard LIST POSITIONS, -1 where the reference returns None."""
        import torch
        from . import ops
        kb = keys if isinstance(keys, ops.KeyBatch) else ops.KeyBatch.from_strings(keys)
        mv = self.registry.get_model_version(model_name, version)
        if mv is None or not mv.shards:
            return torch.full((len(kb),), -1, dtype=torch.int32, device=kb.key_off.device)
        routable, healthy = self.health_masks(mv.shards, prefer_healthy)
        dev = kb.key_off.device
        r = torch.tensor(routable, dtype=torch.uint8, device=dev)
        h = torch.tensor(healthy, dtype=torch.int32, device=dev) if healthy else None
        shard, _, _ = ops.route_md5(kb, len(mv.shards), r, h, self.failover_enabled, want_digests=False)
        return shard

    def route_request(self, model_name: str, version: str, request_key: str,
                      prefer_healthy: bool = True) -> Optional[ModelShard]:
        mv = self.registry.get_model_version(model_name, version)
        if mv is None or not mv.shards:
            return None
        pos = int(self.route(model_name, version, [request_key], prefer_healthy)[0])
        return None if pos < 0 else mv.shards[pos]

    # -- control plane: health probe loop (src/router.py:247-306), kept host-side -------------------
    async def start(self) -> None:
        if self._running:
            return
        self._running = True
        self._health_check_task = asyncio.create_task(self._health_check_loop())

    async def stop(self) -> None:
        self._running = False
        if self._health_check_task:
            self._health_check_task.cancel()
            try:
                await self._health_check_task
            except asyncio.CancelledError:
                pass

    async def _health_check_loop(self) -> None:
        while self._running:
            try:
                await asyncio.sleep(self.health_check_interval)
                await asyncio.gather(*[self._check_worker_health(w) for w in list(self.workers)],
                                     return_exceptions=True)
            except asyncio.CancelledError:
                break

    async def _check_worker_health(self, worker_id: str) -> None:
        w = self.workers.get(worker_id)
        if w is None:
            return
        try:
            host, port = w.address.split(":")
            _, writer = await asyncio.wait_for(asyncio.open_connection(host, int(port)),
                                               timeout=self.health_check_timeout)
            writer.close()
            await writer.wait_closed()
            self.mark_worker_success(worker_id)
        except Exception:  # noqa: BLE001 - any failure to connect counts against the worker
            self.mark_worker_failure(worker_id)
        w.last_health_check = time.time()

    # -- introspection (src/router.py:308-358) ----------------------------------------------------------
    def get_worker_address(self, worker_id: str) -> Optional[str]:
        w = self.workers.get(worker_id)
        return w.address if w else None

    def get_healthy_workers(self) -> List[str]:
        return [i for i, w in self.workers.items() if w.health == WorkerHealth.HEALTHY]

    def get_unhealthy_workers(self) -> List[str]:
        return [i for i, w in self.workers.items() if w.health == WorkerHealth.UNHEALTHY]

    def get_stats(self) -> Dict[str, Any]:
        total, ok, bad = len(self.workers), len(self.get_healthy_workers()), len(self.get_unhealthy_workers())
        return {"total_workers": total, "healthy_workers": ok, "unhealthy_workers": bad,
                "unknown_workers": total - ok - bad, "failover_enabled": self.failover_enabled,
                "health_check_interval": self.health_check_interval}

    def get_worker_info(self, worker_id: str) -> Optional[Dict[str, Any]]:
        w = self.workers.get(worker_id)
        if w is None:
            return None
        return {"worker_id": w.worker_id, "address": w.address, "health": w.health.name,
                "consecutive_failures": w.consecutive_failures, "last_health_check": w.last_health_check,
                "last_success": w.last_success, "metadata": w.metadata}
