"""Composed CPU oracle for the hot path (TEST INFRASTRUCTURE).

The reference never chains its four components (README.md:56-60 describes a
coordinator.py that does not exist), so the composition is the one
BASELINE.json's north_star names and SURVEY.md call-stack (E) spells out:

  route_request per key (router.py:140)  -> group by owner shard, request order kept
  owner: KVCache.get per key, in order   (kvstore.py:136)
         misses: PACK   inputs_list = [req.inputs ...]      (batcher.py:211)
                 INFER  results = callback(model, version, inputs_list) (batcher.py:212)
                 UNPACK result i -> request i                (batcher.py:225-227)
         KVCache.set(key, result) for each miss, in order    (kvstore.py:104)
  hits return the cached result; unroutable requests (route -> None) return None.

With several ingest ranks the owner applies the sources in rank order
(SURVEY.md section 8(e)); ``process_exchange`` does exactly that.
"""
from __future__ import annotations

import asyncio
from typing import Any, Callable, List, Optional, Sequence

from . import ref_port as rp


class OraclePipeline:
    def __init__(self, n_shards: int, max_size: int, default_ttl: Optional[float],
                 model: Callable[[List[Any]], List[Any]], model_name: str = "m", version: str = "1",
                 failover_enabled: bool = True):
        self.now = 0.0
        self.model_name, self.version = model_name, version
        self.registry = rp.PortRegistry()
        self.registry.register_model(model_name, version)
        self.router = rp.PortRouter(self.registry, failover_enabled=failover_enabled)
        for s in range(n_shards):
            self.registry.add_shard(model_name, version, s, f"w{s}")
            self.router.register_worker(f"w{s}")
            self.router.mark_worker_success(f"w{s}")
        self.caches = [rp.PortKVCache(max_size, default_ttl, clock=lambda: self.now)
                       for _ in range(n_shards)]
        self.model = model
        self.n_shards = n_shards

    def route(self, keys: Sequence[str], prefer_healthy: bool = True) -> List[int]:
        shards = self.registry.shards(self.model_name, self.version)
        out = []
        for k in keys:
            s = self.router.route(self.model_name, self.version, k, prefer_healthy)
            out.append(-1 if s is None else shards.index(s))
        return out

    def _owner(self, s: int, keys, payloads, idx, hit, out) -> None:
        cache = self.caches[s]
        miss = []
        for i in idx:
            v = cache.get(keys[i])
            if v is None:
                miss.append(i)
            else:
                hit[i] = 1
                out[i] = v
        inputs_list = [payloads[i] for i in miss]            # PACK
        results = self.model(inputs_list) if miss else []    # INFER
        if len(results) != len(miss):
            raise ValueError(f"Batch callback returned {len(results)} results but expected {len(miss)}")
        for i, r in zip(miss, results):                      # UNPACK
            out[i] = r
        for i in miss:
            cache.set(keys[i], out[i])

    def process(self, keys: Sequence[str], payloads: Sequence[Any], now: float,
                prefer_healthy: bool = True) -> dict:
        """One flushed batch ingested on one rank."""
        return self.process_exchange([(keys, payloads)], now, prefer_healthy)[0]

    def process_exchange(self, per_rank, now: float, prefer_healthy: bool = True) -> List[dict]:
        """``per_rank`` = [(keys, payloads), ...] one tuple per ingest rank.  Owners see the
        sources concatenated in rank order, request order kept inside each source."""
        self.now = now
        flat_keys, flat_pay, spans = [], [], []
        for keys, pays in per_rank:
            spans.append((len(flat_keys), len(keys)))
            flat_keys.extend(keys)
            flat_pay.extend(pays)
        shard = self.route(flat_keys, prefer_healthy)
        n = len(flat_keys)
        hit, out = [0] * n, [None] * n
        for s in range(self.n_shards):
            idx = [i for i in range(n) if shard[i] == s]
            if idx:
                self._owner(s, flat_keys, flat_pay, idx, hit, out)
        return [{"shard": shard[a:a + m], "hit": hit[a:a + m], "out": out[a:a + m]} for a, m in spans]

    def state(self) -> List[dict]:
        return [{"lru": c.lru_order(), **c.stats()} for c in self.caches]


class BaselinePipeline(OraclePipeline):
    """Same composition, but the pack/infer/unpack leg runs through the asyncio batcher
    restatement (add_request per miss -> size/stop flush -> futures), i.e. the path the
    reference's own classes would take.  Used for the timed CPU baseline only."""

    def __init__(self, *a, max_batch_size: int = 4096, **kw):
        super().__init__(*a, **kw)
        self.max_batch_size = max_batch_size
        self.loop = asyncio.new_event_loop()

    def _owner(self, s, keys, payloads, idx, hit, out) -> None:
        cache = self.caches[s]
        miss = []
        for i in idx:
            v = cache.get(keys[i])
            if v is None:
                miss.append(i)
            else:
                hit[i] = 1
                out[i] = v

        async def cb(model, version, inputs_list):
            return self.model(inputs_list)

        async def run():
            b = rp.PortBatcher(self.max_batch_size, 1000.0, cb)
            await b.start()
            futs = [await b.add_request(self.model_name, self.version, payloads[i]) for i in miss]
            await b.stop()
            return [await f for f in futs]

        for i, r in zip(miss, self.loop.run_until_complete(run()) if miss else []):
            out[i] = r
        for i in miss:
            cache.set(keys[i], out[i])
