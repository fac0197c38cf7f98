"""The reference's OWN classes composed into the hot path (TEST INFRASTRUCTURE; timed CPU baseline only).

`oracle/_ref/src` is a verbatim copy of the reference package (oracle/fetch_ref.py).  This module imports
`src.router.Router`, `src.model_registry.ModelRegistry`, `src.kvstore.KVCache` and `src.batcher.Batcher` from it,
unmodified, and chains them the way SURVEY.md call-stack (E) and tests/golden/make_golden.py::gen_pipeline do:

    Router.route_request per key              (src/router.py:140)
    owner: KVCache.get per key, in order      (src/kvstore.py:136)
           misses -> Batcher.add_request each (src/batcher.py:102; size flush -> _process_batch :202:
                     PACK :211, callback = the mock model, UNPACK :225-227)
           KVCache.set(key, result) per miss  (src/kvstore.py:104)

Two interventions, the ones SURVEY.md section 8(c) lists: the cache's clock is injected (`src.kvstore.time`)
and the mock model's sleep is left out (the callback is the model function itself).  Logging is silenced.
"""
from __future__ import annotations

import asyncio
import logging
import os
import sys
import types
from typing import Any, Callable, List, Optional, Sequence

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.path.join(HERE, "_ref")


def available() -> bool:
    return os.path.isfile(os.path.join(REF_ROOT, "src", "kvstore.py"))


class ReferencePipeline:
    """Same surface as oracle.pipeline.BaselinePipeline.process(), running the reference's classes."""

    def __init__(self, n_shards: int, max_size: int, default_ttl: Optional[float],
                 model: Callable[[List[Any]], List[Any]], max_batch_size: int = 4096):
        if not available():
            raise ImportError("oracle/_ref/src is missing: run oracle/fetch_ref.py where /root/reference exists")
        if REF_ROOT not in sys.path:
            sys.path.insert(0, REF_ROOT)
        sys.dont_write_bytecode = True
        logging.disable(logging.CRITICAL)
        import src.kvstore as ref_kv
        from src.batcher import Batcher
        from src.model_registry import ModelRegistry
        from src.router import Router
        self.now = 0.0
        ref_kv.time = types.SimpleNamespace(time=lambda: self.now)   # injected clock (kvstore.py:8 `import time`)
        self._Batcher = Batcher
        self.registry = ModelRegistry()
        self.registry.register_model("m", "1", "/p", {}, {})
        self.router = Router(self.registry)
        for s in range(n_shards):
            self.registry.add_shard("m", "1", s, f"w{s}")
            self.router.register_worker(f"w{s}", f"127.0.0.1:{9000 + s}")
            self.router.mark_worker_success(f"w{s}")
        self.shards = self.registry.get_model_version("m", "1").shards
        self.caches = [ref_kv.KVCache(max_size=max_size, eviction_policy="lru", default_ttl=default_ttl)
                       for _ in range(n_shards)]
        self.model, self.max_batch_size = model, max_batch_size
        self.loop = asyncio.new_event_loop()

    async def _infer(self, inputs_list):
        async def cb(model, version, rows):
            return self.model(rows)
        b = self._Batcher(max_batch_size=self.max_batch_size, max_latency_ms=1000.0, batch_callback=cb)
        await b.start()
        futs = [await b.add_request("m", "1", x) for x in inputs_list]
        await b.stop()
        return [await f for f in futs]

    def process(self, keys: Sequence[str], payloads: Sequence[Any], now: float) -> dict:
        self.now = now
        n = len(keys)
        shard = []
        for k in keys:
            sh = self.router.route_request("m", "1", k)
            shard.append(-1 if sh is None else self.shards.index(sh))
        hit, out = [0] * n, [None] * n
        for s in range(len(self.caches)):
            cache = self.caches[s]
            idx = [i for i in range(n) if shard[i] == s] if len(self.caches) > 1 else range(n)
            miss = []
            for i in idx:
                v = cache.get(keys[i])
                if v is None:
                    miss.append(i)
                else:
                    hit[i], out[i] = 1, v
            if miss:
                for i, r in zip(miss, self.loop.run_until_complete(self._infer([payloads[i] for i in miss]))):
                    out[i] = r
                for i in miss:
                    cache.set(keys[i], out[i])
        return {"shard": shard, "hit": hit, "out": out}

    def state(self) -> List[dict]:
        return [{"lru": list(c._access_order), **c.get_stats()} for c in self.caches]
