"""Drop-in mirror of the reference's LoadBalancer (src/load_balancer.py): the alternative, sequential worker picker.

Control-plane logic, kept in Python as SURVEY.md section 8(f4) says: four strategies over the workers whose
consecutive-failure count is below `max_failures` (round robin, fewest active connections, random, lowest average
latency), per-worker statistics, and a background health check.  Same names, arguments, return values and quirks as
the reference (load_balancer.py:18-345) -- round robin advances its cursor before picking (:243-244), a pinned
`worker_id` that is unhealthy yields None instead of falling back (:144-147), `min()` ties go to the first healthy
worker in registration order.  The health probe is pluggable; by default it opens and closes a TCP connection like the
reference (:315-345)."""
from __future__ import annotations

import asyncio
import logging
import random
import time
from dataclasses import dataclass, field
from enum import Enum
from typing import Any, Awaitable, Callable, Dict, List, Optional, Tuple

logger = logging.getLogger(__name__)


class LoadBalancerStrategy(Enum):
    ROUND_ROBIN = "round_robin"
    LEAST_CONNECTIONS = "least_connections"
    RANDOM = "random"
    LEAST_LATENCY = "least_latency"


@dataclass
class WorkerStats:
    request_count: int = 0
    error_count: int = 0
    active_connections: int = 0
    total_latency: float = 0.0
    last_seen: float = field(default_factory=time.time)

    @property
    def avg_latency(self) -> float:
        return self.total_latency / self.request_count if self.request_count > 0 else 0.0


async def _tcp_probe(address: str, timeout: float) -> None:
    host, port = address.split(":")
    _reader, writer = await asyncio.wait_for(asyncio.open_connection(host, int(port)), timeout=timeout)
    writer.close()
    await writer.wait_closed()


class LoadBalancer:
    def __init__(self, strategy: LoadBalancerStrategy = LoadBalancerStrategy.ROUND_ROBIN, health_check_interval: float = 5.0,
                 health_check_timeout: float = 2.0, max_failures: int = 3,
                 probe: Optional[Callable[[str, float], Awaitable[None]]] = None):
        self.strategy = strategy
        self.health_check_interval = health_check_interval
        self.health_check_timeout = health_check_timeout
        self.max_failures = max_failures
        self.workers: Dict[str, str] = {}
        self.worker_stats: Dict[str, WorkerStats] = {}
        self.health_checks: Dict[str, int] = {}
        self._pickers = {LoadBalancerStrategy.ROUND_ROBIN: self._round_robin,
                         LoadBalancerStrategy.LEAST_CONNECTIONS: self._least_connections,
                         LoadBalancerStrategy.RANDOM: self._random,
                         LoadBalancerStrategy.LEAST_LATENCY: self._least_latency}
        self._current_worker_idx = 0
        self._running = False
        self._health_check_task: Optional[asyncio.Task] = None
        self._probe = probe or _tcp_probe

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

    def register_worker(self, worker_id: str, address: str) -> None:
        self.workers[worker_id] = address
        if worker_id not in self.worker_stats:   # re-registering keeps the history (load_balancer.py:105-107)
            self.worker_stats[worker_id] = WorkerStats()
            self.health_checks[worker_id] = 0

    def unregister_worker(self, worker_id: str) -> bool:
        if worker_id not in self.workers:
            return False
        del self.workers[worker_id]
        del self.worker_stats[worker_id]
        self.health_checks.pop(worker_id, None)
        return True

    def _healthy(self, worker_id: str) -> bool:
        return self.health_checks.get(worker_id, 0) < self.max_failures

    async def get_worker(self, worker_id: Optional[str] = None) -> Optional[Tuple[str, str]]:
        if not self.workers:
            return None
        if worker_id and worker_id in self.workers:   # pinned: that worker or nothing
            return (worker_id, self.workers[worker_id]) if self._healthy(worker_id) else None
        candidates = [wid for wid in self.workers if self._healthy(wid)]
        if not candidates:
            return None
        picked = await self._pickers.get(self.strategy, self._round_robin)(candidates)
        return (picked, self.workers[picked]) if picked else None

    async def update_stats(self, worker_id: str, success: bool, latency: float) -> None:
        stats = self.worker_stats.setdefault(worker_id, WorkerStats())
        stats.request_count += 1
        stats.total_latency += latency
        stats.last_seen = time.time()
        if success:
            self.health_checks[worker_id] = 0
        else:
            stats.error_count += 1
            self.health_checks[worker_id] = self.health_checks.get(worker_id, 0) + 1

    def get_worker_stats(self, worker_id: str) -> Optional[Dict[str, Any]]:
        stats = self.worker_stats.get(worker_id)
        if stats is None:
            return None
        return {"worker_id": worker_id, "address": self.workers.get(worker_id, "unknown"),
                "request_count": stats.request_count, "error_count": stats.error_count,
                "active_connections": stats.active_connections, "avg_latency": stats.avg_latency,
                "last_seen": stats.last_seen, "healthy": self._healthy(worker_id)}

    def get_all_stats(self) -> Dict[str, Dict[str, Any]]:
        return {wid: self.get_worker_stats(wid) for wid in self.workers}

    def __repr__(self) -> str:
        return f"LoadBalancer(workers={self.workers})"

    async def _round_robin(self, worker_ids: List[str]) -> Optional[str]:
        if not worker_ids:
            return None
        self._current_worker_idx = (self._current_worker_idx + 1) % len(worker_ids)
        return worker_ids[self._current_worker_idx]

    async def _least_connections(self, worker_ids: List[str]) -> Optional[str]:
        return min(worker_ids, key=lambda wid: self.worker_stats[wid].active_connections) if worker_ids else None

    async def _random(self, worker_ids: List[str]) -> Optional[str]:
        return random.choice(worker_ids) if worker_ids else None

    async def _least_latency(self, worker_ids: List[str]) -> Optional[str]:
        return min(worker_ids, key=lambda wid: self.worker_stats[wid].avg_latency) if worker_ids else None

    async def _health_check_loop(self) -> None:
        while self._running:
            try:
                await asyncio.sleep(self.health_check_interval)
                await self._check_all_workers()
            except asyncio.CancelledError:
                break
            except Exception as e:  # noqa: BLE001
                logger.error(f"Error in health check loop: {e}")

    async def _check_all_workers(self) -> None:
        if self.workers:
            await asyncio.gather(*[self._check_worker(wid) for wid in list(self.workers)], return_exceptions=True)

    async def _check_worker(self, worker_id: str) -> None:
        if worker_id not in self.workers:
            return
        try:
            t0 = time.time()
            await self._probe(self.workers[worker_id], self.health_check_timeout)
            latency = time.time() - t0
            self.health_checks[worker_id] = 0
            stats = self.worker_stats[worker_id]
            stats.last_seen = time.time()
            stats.total_latency += latency
            stats.request_count += 1
        except (asyncio.TimeoutError, ConnectionError) as e:
            logger.warning(f"Health check failed for worker {worker_id}: {e}")
            self.health_checks[worker_id] = self.health_checks.get(worker_id, 0) + 1
