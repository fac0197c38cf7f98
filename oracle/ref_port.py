"""Pure-Python restatement of the reference's hot-path algorithms (TEST INFRASTRUCTURE).

Each class/function cites the reference ``file:line`` it follows (paths relative
to the reference checkout).  The port keeps the reference's data structures
(``hashlib.md5``, ``dict`` + ``OrderedDict``) so that timing it is a fair stand-in
for the reference's CPU path, but the clock is injected (``clock()``) instead of
``time.time()`` so traces are reproducible, and sleeps are optional.

Checked against the reference itself through the fixtures in ``tests/golden``
(see ``tests/golden/make_golden.py`` and ``tests/test_oracle_golden.py``).
"""
from __future__ import annotations

import asyncio
import hashlib
import uuid
from collections import OrderedDict
from typing import Any, Awaitable, Callable, Dict, List, Optional, Sequence

# ----------------------------------------------------------------------------
# Routing  (src/model_registry.py:149-161, src/router.py:140-221)
# ----------------------------------------------------------------------------

HEALTHY, UNHEALTHY, UNKNOWN = "HEALTHY", "UNHEALTHY", "UNKNOWN"


def md5_digest(key: str) -> bytes:
    """16 digest bytes of the UTF-8 key (model_registry.py:159, ``key.encode()``)."""
    return hashlib.md5(key.encode()).digest()


def key_hash(key: str) -> int:
    """``int(md5(key).hexdigest(), 16)`` == big-endian integer of the digest
    (model_registry.py:159, router.py:214)."""
    return int.from_bytes(md5_digest(key), "big")


def shard_index(key: str, n_shards: int) -> int:
    """List position of the shard that owns ``key`` (model_registry.py:160)."""
    return key_hash(key) % n_shards


class PortRegistry:
    """Shard lists per (model, version); restates ModelRegistry.add_shard /
    get_shard_for_key (model_registry.py:116-161).  A shard is ``(shard_id, worker_id)``."""

    def __init__(self) -> None:
        self.versions: Dict[tuple, List[tuple]] = {}

    def register_model(self, model: str, version: str) -> None:
        # model_registry.py:98-114 always installs a fresh ModelVersion (shards reset)
        self.versions[(model, version)] = []

    def add_shard(self, model: str, version: str, shard_id: int, worker_id: str) -> tuple:
        if (model, version) not in self.versions:
            raise ValueError(f"Model {model} version {version} not found")  # :125-126
        shards = self.versions[(model, version)]
        for s in shards:  # idempotent on shard_id (:129-132)
            if s[0] == shard_id:
                return s
        s = (shard_id, worker_id)
        shards.append(s)
        return s

    def shards(self, model: str, version: str) -> Optional[List[tuple]]:
        return self.versions.get((model, version))

    def shard_for_key(self, model: str, version: str, key: str) -> Optional[tuple]:
        shards = self.versions.get((model, version))
        if not shards:  # unknown model/version or empty shard list (:151-156)
            return None
        return shards[shard_index(key, len(shards))]


class PortRouter:
    """Restates Router.route_request + _find_alternative_shard + health marks
    (router.py:140-245).  ``workers`` maps worker_id -> [health, consecutive_failures]."""

    def __init__(self, registry: PortRegistry, max_consecutive_failures: int = 3,
                 failover_enabled: bool = True) -> None:
        self.registry = registry
        self.max_consecutive_failures = max_consecutive_failures
        self.failover_enabled = failover_enabled
        self.workers: Dict[str, list] = {}

    def register_worker(self, worker_id: str) -> None:
        self.workers[worker_id] = [UNKNOWN, 0]  # router.py:127-132 (new workers start UNKNOWN)

    def unregister_worker(self, worker_id: str) -> None:
        self.workers.pop(worker_id, None)

    def mark_worker_success(self, worker_id: str) -> None:  # router.py:223-231
        w = self.workers.get(worker_id)
        if w is not None:
            w[1] = 0
            w[0] = HEALTHY

    def mark_worker_failure(self, worker_id: str) -> None:  # router.py:233-245
        w = self.workers.get(worker_id)
        if w is not None:
            w[1] += 1
            if w[1] >= self.max_consecutive_failures:
                w[0] = UNHEALTHY

    def _alternative(self, model: str, version: str, key: str) -> Optional[tuple]:
        shards = self.registry.shards(model, version)
        if shards is None:
            return None
        healthy = [s for s in shards
                   if s[1] in self.workers and self.workers[s[1]][0] == HEALTHY]  # :197-201
        if not healthy:
            return None
        return healthy[key_hash(key) % len(healthy)]  # :214-215

    def route(self, model: str, version: str, key: str, prefer_healthy: bool = True) -> Optional[tuple]:
        shard = self.registry.shard_for_key(model, version, key)
        if shard is None:
            return None
        w = self.workers.get(shard[1])
        if w is None:  # :167-173
            return self._alternative(model, version, key) if self.failover_enabled else None
        if prefer_healthy and w[0] != HEALTHY:  # :175-182
            return self._alternative(model, version, key) if self.failover_enabled else None
        return shard


# ----------------------------------------------------------------------------
# KV cache, LRU policy + TTL  (src/kvstore.py)
# ----------------------------------------------------------------------------

class PortEntry:
    __slots__ = ("value", "created_at", "last_accessed", "access_count", "ttl")

    def __init__(self, value, now, ttl):
        self.value = value
        self.created_at = now
        self.last_accessed = now
        self.access_count = 0
        self.ttl = ttl


class PortKVCache:
    """Restates KVCache (kvstore.py:26-236) for the three eviction policies: "lru" (OrderedDict of
    accesses), "lfu" (dict key -> use count; victim = first key with the smallest count, kvstore.py:94-98) and
    "fifo" (victim = first key of the cache dict itself, kvstore.py:99-102).

    ``clock`` replaces ``time.time`` (kvstore.py:78,127,128,159,204)."""

    def __init__(self, max_size: int = 1000, default_ttl: Optional[float] = None,
                 clock: Callable[[], float] = None, policy: str = "lru") -> None:
        self.max_size = max_size
        self.default_ttl = default_ttl
        self.clock = clock
        self.policy = policy.lower()
        self.cache: Dict[str, PortEntry] = {}
        self.order: "OrderedDict[str, float]" = OrderedDict()  # LRU list, head = oldest
        self.counts: Dict[str, int] = {}                        # LFU: _access_count (kvstore.py:67-68)
        self.hits = self.misses = self.evictions = 0

    # kvstore.py:200-204 -- strict '>' ; ttl None never expires
    def _expired(self, e: PortEntry) -> bool:
        return e.ttl is not None and (self.clock() - e.created_at) > e.ttl

    # kvstore.py:71-78
    def _touch(self, key: str) -> None:
        if self.policy == "lru":
            if key in self.order:
                self.order.move_to_end(key)
            else:
                self.order[key] = self.clock()
        elif self.policy == "lfu":  # kvstore.py:79-80
            self.counts[key] = self.counts.get(key, 0) + 1

    # kvstore.py:82-93 -- counter bumped BEFORE the pop, so it counts even if the pop raises
    def _make_room(self) -> None:
        if len(self.cache) < self.max_size:
            return
        self.evictions += 1
        if self.policy == "lru":
            victim = next(iter(self.order))  # StopIteration when nothing is stamped
            del self.order[victim]
        elif self.policy == "lfu":
            victim = min(self.counts.items(), key=lambda x: x[1])[0]  # ValueError when nothing is counted
            del self.counts[victim]
        else:
            victim = next(iter(self.cache))  # FIFO: insertion order of the cache dict; StopIteration when empty
        del self.cache[victim]

    # kvstore.py:104-134
    def set(self, key: str, value: Any, ttl: Optional[float] = None, batch: bool = False) -> None:
        if key in self.cache:
            self.delete(key)
        self._make_room()
        now = self.clock()
        self.cache[key] = PortEntry(value, now, ttl if ttl is not None else self.default_ttl)
        if not batch:
            self._touch(key)

    # kvstore.py:136-164
    def get(self, key: str, default: Any = None) -> Any:
        e = self.cache.get(key)
        if e is None:
            self.misses += 1
            return default
        if self._expired(e):
            self.delete(key)
            self.misses += 1
            return default
        e.last_accessed = self.clock()
        e.access_count += 1
        self._touch(key)
        self.hits += 1
        return e.value

    def batch_get(self, keys: Sequence[str]) -> Dict[str, Any]:  # kvstore.py:166-168
        return {k: self.get(k) for k in keys}

    def batch_set(self, items: Dict[str, Any], ttl: Optional[float] = None) -> None:  # :170-176
        for k, v in items.items():
            self.set(k, v, ttl=ttl, batch=True)
        for k in items:
            self._touch(k)

    def delete(self, key: str) -> bool:  # kvstore.py:178-187
        if key in self.cache:
            del self.cache[key]
            self.order.pop(key, None)
            self.counts.pop(key, None)
            return True
        return False

    def clear(self) -> None:  # kvstore.py:189-198
        self.cache.clear()
        self.order.clear()
        self.counts.clear()
        self.hits = self.misses = self.evictions = 0

    def contains(self, key: str) -> bool:  # kvstore.py:221-228
        e = self.cache.get(key)
        if e is None:
            return False
        if self._expired(e):
            self.delete(key)
            return False
        return True

    def length(self) -> int:  # kvstore.py:230-236 (purges every expired entry first)
        for k in [k for k, e in self.cache.items() if self._expired(e)]:
            self.delete(k)
        return len(self.cache)

    def stats(self) -> Dict[str, Any]:  # kvstore.py:206-219
        total = self.hits + self.misses
        return {"hits": self.hits, "misses": self.misses,
                "hit_rate": self.hits / total if total > 0 else 0,
                "size": len(self.cache), "max_size": self.max_size,
                "evictions": self.evictions, "eviction_policy": self.policy}

    def lru_order(self) -> List[str]:
        return list(self.order)


# ----------------------------------------------------------------------------
# Batcher  (src/batcher.py) -- asyncio restatement of the order / error contract
# ----------------------------------------------------------------------------

class PortBatcher:
    """Restates Batcher.add_request / _flush_batch / _schedule_flush / _process_batch
    (batcher.py:102-240): per "model:version" buffers, inline size flush, latency
    flush task, pack = gather inputs in arrival order (:211), unpack = result i ->
    request i (:225-227), any failure fans out to every future of the batch (:235-240)."""

    def __init__(self, max_batch_size: int = 32, max_latency_ms: float = 100.0,
                 batch_callback: Optional[Callable[[str, str, List[Any]], Awaitable[List[Any]]]] = None):
        if max_batch_size < 1:
            raise ValueError("max_batch_size must be at least 1")
        if max_latency_ms <= 0:
            raise ValueError("max_latency_ms must be greater than 0")
        self.max_batch_size = max_batch_size
        self.max_latency = max_latency_ms / 1000.0
        self.batch_callback = batch_callback
        self.pending: Dict[str, dict] = {}
        self.timers: Dict[str, asyncio.Task] = {}
        self.running = False
        self.lock = asyncio.Lock()
        self.total_batches = self.total_requests = self.total_batched_requests = 0

    async def start(self) -> None:
        self.running = True

    async def stop(self) -> None:  # batcher.py:71-100
        if not self.running:
            return
        self.running = False
        async with self.lock:
            timers = list(self.timers.values())
            left = list(self.pending.items())
            self.timers.clear()
            self.pending.clear()
        for t in timers:
            if not t.done():
                t.cancel()
                try:
                    await t
                except asyncio.CancelledError:
                    pass
        for key, b in left:
            if b["reqs"]:
                await self._process(key, b, list(b["reqs"]))

    async def add_request(self, model: str, version: str, inputs: Any,
                          request_id: Optional[str] = None) -> asyncio.Future:
        if not self.running:
            raise RuntimeError("Batcher is not running")
        if request_id is None:
            request_id = str(uuid.uuid4())
        key = f"{model}:{version}"
        req = {"id": request_id, "inputs": inputs, "future": asyncio.get_running_loop().create_future()}
        async with self.lock:
            if key not in self.pending:
                b = {"model": model, "version": version, "reqs": []}
                self.pending[key] = b
                self.timers[key] = asyncio.create_task(self._latency_flush(key, b))
            b = self.pending[key]
            b["reqs"].append(req)
            self.total_requests += 1
            full = len(b["reqs"]) >= self.max_batch_size
        if full:
            await self._size_flush(key, b)
        return req["future"]

    async def _latency_flush(self, key: str, b: dict) -> None:  # batcher.py:151-178
        try:
            await asyncio.sleep(self.max_latency)
            async with self.lock:
                if self.pending.get(key) is b and b["reqs"]:
                    reqs = list(b["reqs"])
                    b["reqs"].clear()
                    del self.pending[key]
                    await self._process(key, b, reqs)
        except asyncio.CancelledError:
            pass

    async def _size_flush(self, key: str, b: dict) -> None:  # batcher.py:180-200
        async with self.lock:
            if not b["reqs"]:
                return
            reqs = list(b["reqs"])
            b["reqs"].clear()
            if self.pending.get(key) is b:
                del self.pending[key]
            t = self.timers.pop(key, None)
            if t is not None and not t.done():
                t.cancel()
        await self._process(key, b, reqs)

    async def _process(self, key: str, b: dict, reqs: List[dict]) -> None:  # batcher.py:202-240
        if not reqs:
            return
        self.total_batches += 1
        self.total_batched_requests += len(reqs)
        try:
            if self.batch_callback:
                inputs_list = [r["inputs"] for r in reqs]  # PACK (:211)
                results = await self.batch_callback(b["model"], b["version"], inputs_list)
                if len(results) != len(reqs):
                    raise ValueError(f"Batch callback returned {len(results)} results "
                                     f"but expected {len(reqs)}")
                for r, res in zip(reqs, results):  # UNPACK (:225-227)
                    if not r["future"].done():
                        r["future"].set_result(res)
            else:
                err = RuntimeError("No batch callback configured")
                for r in reqs:
                    if not r["future"].done():
                        r["future"].set_exception(err)
        except Exception as e:  # noqa: BLE001 - mirrors the reference's blanket fan-out
            for r in reqs:
                if not r["future"].done():
                    r["future"].set_exception(e)

    async def get_stats(self) -> Dict[str, Any]:  # batcher.py:242-269 (duplicate key: last wins)
        avg = self.total_batched_requests / self.total_batches if self.total_batches > 0 else 0
        async with self.lock:
            pend_req = sum(len(b["reqs"]) for b in self.pending.values() if b["reqs"])
        return {"total_batches": self.total_batches, "total_requests": self.total_requests,
                "total_batched_requests": self.total_batched_requests,
                "pending_batches": len(self.pending), "pending_requests": pend_req,
                "avg_batch_size": avg, "max_batch_size": self.max_batch_size,
                "max_latency_ms": self.max_latency * 1000.0}


# ----------------------------------------------------------------------------
# Mock models and worker  (src/mock_models/*.py, src/worker.py:135-184)
# ----------------------------------------------------------------------------

async def port_mock_batch_inference(model: str, version: str, inputs: List[Any], **kw) -> List[dict]:
    """Restates mock_batch_inference (mock_inference.py:12-73); ``timestamp`` omitted."""
    latency_ms = kw.get("latency_ms", 100)
    error_rate = kw.get("error_rate", 0.0)
    n = len(inputs)
    if latency_ms:
        await asyncio.sleep(latency_ms / 1000.0)
    out = []
    for i, x in enumerate(inputs):
        if error_rate > 0 and (i / n) < error_rate:  # :46
            out.append({"success": False, "error": f"Simulated error processing input {i}",
                        "model": model, "version": version, "input_id": i})
        else:
            out.append({"success": True, "result": f"Processed by {model}:{version} - {x}",
                        "model": model, "version": version, "input_id": i,
                        "metadata": {"processing_time_ms": latency_ms, "batch_size": n,
                                     "batch_position": i}})
    return out


class PortTickBatcher:
    """The Batcher's flush schedule restated for time that advances in ticks (the form csrc/batcher.cu implements).

    The reference arms a timer of max_latency when a key's batch is opened (batcher.py:124-135), flushes the batch
    as soon as it holds max_batch_size requests (:143-146) and flushes whatever is there when the timer fires
    (:150-166); stop() flushes the open batches in creation order (:70-100).  With no arrivals between two ticks,
    firing the due timers at the start of a tick (in the order they were armed) gives the same batches in the same
    order.  tests/golden/make_golden.py::gen_tick_batcher checks exactly that against the reference's own class run
    on a virtual-time event loop.  Returns flushed batches as (key, [inputs...], trigger)."""

    def __init__(self, max_batch_size: int = 32, max_latency_ms: float = 50.0):
        if max_batch_size <= 0:
            raise ValueError("max_batch_size must be positive")
        if max_latency_ms <= 0:
            raise ValueError("max_latency_ms must be positive")
        self.max_batch_size = max_batch_size
        self.max_latency = max_latency_ms / 1000.0
        self.open: Dict[str, dict] = {}  # insertion order = creation order (batcher.py:88 iterates it)
        self.total_batches = self.total_requests = self.total_batched_requests = 0

    def _flush(self, key: str, trigger: str, out: list) -> None:
        b = self.open.pop(key)
        self.total_batches += 1                                  # batcher.py:206
        self.total_batched_requests += len(b["requests"])        # :207
        out.append((key, list(b["requests"]), trigger))          # PACK: arrival order (:211)

    def tick(self, now: float, arrivals, force: bool = False) -> list:
        """arrivals: [(key, inputs)] in arrival order."""
        out: list = []
        for key in [k for k, b in self.open.items() if now - b["created_at"] >= self.max_latency]:
            self._flush(key, "latency", out)                     # _schedule_flush (:150-166), oldest timer first
        for key, inputs in arrivals:
            if key not in self.open:
                self.open[key] = {"created_at": now, "requests": []}   # :124-135
            self.open[key]["requests"].append(inputs)
            self.total_requests += 1                             # :141
            if len(self.open[key]["requests"]) >= self.max_batch_size:  # :143
                self._flush(key, "size", out)
        if force:                                                # stop() (:84-97)
            for key in list(self.open):
                self._flush(key, "forced", out)
        return out

    def pending(self) -> Dict[str, int]:
        return {k: len(b["requests"]) for k, b in self.open.items()}


class PortFakeModel:
    """Restates FakeModel.predict (fake_model.py:33-67): an echo.  ``sleep`` False
    removes the 50-150 ms simulated latency; timestamp/request_id omitted."""

    def __init__(self, model_name: str, sleep: bool = False, clock: Callable[[], float] = None):
        self.model_name = model_name
        self.sleep = sleep
        self.clock = clock
        self.request_count = 0
        self.error_count = 0

    async def predict(self, inputs: Any) -> dict:
        self.request_count += 1
        if self.sleep:
            await asyncio.sleep(0.05 + (self.clock() % 0.1))  # :47
        return {"model": self.model_name, "output": inputs,
                "metadata": {"batch_size": len(inputs) if isinstance(inputs, (list, tuple)) else 1}}


class PortWorker:
    """Restates Worker._process_request / load_model / unload_model (worker.py:135-184)."""

    def __init__(self, worker_id: str):
        self.worker_id = worker_id
        self.models: Dict[str, PortFakeModel] = {}

    def load_model(self, model_name: str) -> bool:
        if model_name not in self.models:
            self.models[model_name] = PortFakeModel(model_name)
        return True

    def unload_model(self, model_name: str) -> bool:
        return self.models.pop(model_name, None) is not None

    async def process_request(self, request: Any) -> dict:
        if not isinstance(request, dict):
            return {"error": "Request must be a JSON object", "success": False}
        name, inputs = request.get("model"), request.get("inputs")
        if not name or inputs is None:
            return {"error": "Missing required fields: model and inputs are required", "success": False}
        if name not in self.models:
            return {"error": f"Model '{name}' not found", "success": False}
        try:
            outputs = await self.models[name].predict(inputs)
            return {"model": name, "outputs": outputs, "worker_id": self.worker_id, "success": True}
        except Exception as e:  # noqa: BLE001
            return {"error": str(e), "success": False}
