#!/usr/bin/env python
"""Generate the golden fixtures in this directory by EXECUTING THE REFERENCE ITSELF.

Run in the build container only (``/root/reference`` does not exist on the GPU box):

    PYTHONDONTWRITEBYTECODE=1 python tests/golden/make_golden.py

The reference's classes are imported unmodified; the only interventions are the ones
SURVEY.md section 8(c) lists: ``src.kvstore.time`` is replaced by an injected clock,
``asyncio.sleep`` inside ``src.mock_models`` is made instant, logging is silenced and
wall-clock metadata (``timestamp``/``request_id``) is stripped before freezing.
Nothing here is imported by the product or by the tests; the tests read the JSON.
"""
import asyncio
import json
import logging
import os
import random
import sys
import types

REF = os.environ.get("DIE_REFERENCE", "/root/reference")
sys.path.insert(0, REF)
sys.dont_write_bytecode = True
logging.disable(logging.CRITICAL)

import src.kvstore as ref_kv  # noqa: E402
import src.mock_models.fake_model as ref_fake  # noqa: E402
import src.mock_models.mock_inference as ref_mock  # noqa: E402
from src.batcher import Batcher  # noqa: E402
from src.config import ModelConfig  # noqa: E402
from src.model_registry import ModelRegistry  # noqa: E402
from src.router import Router, WorkerHealth  # noqa: E402
from src.worker import Worker  # noqa: E402

HERE = os.path.dirname(os.path.abspath(__file__))
CLOCK = [1000.0]
ref_kv.time = types.SimpleNamespace(time=lambda: CLOCK[0])


async def _no_sleep(_):
    return None


_real_sleep = asyncio.sleep


def dump(name, obj):
    path = os.path.join(HERE, name)
    with open(path, "w") as f:
        json.dump(obj, f, separators=(",", ":"), ensure_ascii=True)
    print(f"{name}: {os.path.getsize(path)} bytes")


# ----------------------------------------------------------------------------- routing
def key_corpus():
    rng = random.Random(7)
    keys = ["user-123", "test_key_1", "test_key_2", "", "a", "session-é", "x" * 55, "x" * 56,
            "x" * 63, "x" * 64, "x" * 65, "x" * 119, "x" * 120, "x" * 200, "中文-key", "\U0001f600",
            "tab\tkey", "nul\x00byte", "message digest", "abc",
            "abcdefghijklmnopqrstuvwxyz",
            "12345678901234567890123456789012345678901234567890123456789012345678901234567890"]
    keys += ["user-%010d" % rng.randrange(10 ** 7) for _ in range(64)]
    alphabet = "abcdefghijklmnopqrstuvwxyzABCDEFGHIJKLMNOPQRSTUVWXYZ0123456789-_:/éü中"
    for n in list(range(0, 70)) + [100, 127, 128, 129, 191, 192, 255, 256, 300, 511, 777]:
        keys.append("".join(rng.choice(alphabet) for _ in range(n)))
    return keys


def make_registry(shard_ids, model="m", version="1"):
    reg = ModelRegistry()
    reg.register_model(model, version, "/p", {}, {})
    for sid in shard_ids:
        reg.add_shard(model, version, sid, f"w{sid}")
    return reg


def gen_route():
    import hashlib
    keys = key_corpus()
    out = {"keys": keys, "md5": [hashlib.md5(k.encode()).hexdigest() for k in keys], "shard_index": {},
           "out_of_order": {}, "failover": []}
    for n in [1, 2, 3, 4, 5, 6, 7, 8, 13, 16, 64, 100, 255, 256, 1000, 65537]:
        reg = make_registry(range(n))
        out["shard_index"][str(n)] = [reg.get_shard_for_key("m", "1", k).shard_id for k in keys]
    ids = [5, 2, 9]
    reg = make_registry(ids)
    out["out_of_order"] = {"shard_ids": ids,
                           "routed_ids": [reg.get_shard_for_key("m", "1", k).shard_id for k in keys]}
    out["unknown_model"] = reg.get_shard_for_key("nope", "1", "k") is None
    reg2 = ModelRegistry()
    reg2.register_model("m", "1", "/p", {}, {})
    out["no_shards"] = reg2.get_shard_for_key("m", "1", "k") is None

    # health / failover decision table over many states
    rng = random.Random(11)
    for case in range(60):
        n = rng.choice([1, 2, 3, 4, 5, 8])
        reg = make_registry(range(n))
        failover = rng.random() < 0.75
        router = Router(reg, failover_enabled=failover, max_consecutive_failures=3)
        states = []
        for s in range(n):
            st = rng.choice(["HEALTHY", "HEALTHY", "UNHEALTHY", "UNKNOWN", "UNREGISTERED"])
            if case == 0:
                st = "UNKNOWN"
            if case == 1:
                st = "HEALTHY"
            states.append(st)
            if st == "UNREGISTERED":
                continue
            router.register_worker(f"w{s}", f"127.0.0.1:{9000 + s}")
            if st == "HEALTHY":
                router.mark_worker_success(f"w{s}")
            elif st == "UNHEALTHY":
                for _ in range(3):
                    router.mark_worker_failure(f"w{s}")
        for prefer in (True, False):
            routed = []
            for k in keys[:48]:
                sh = router.route_request("m", "1", k, prefer_healthy=prefer)
                routed.append(-1 if sh is None else sh.shard_id)
            out["failover"].append({"n": n, "failover_enabled": failover, "states": states,
                                    "prefer_healthy": prefer, "routed": routed})
    # health transitions (router.py:223-245)
    reg = make_registry(range(1))
    router = Router(reg, max_consecutive_failures=3)
    router.register_worker("w0", "127.0.0.1:1")
    trans = [router.workers["w0"].health.name]
    for ev in "ffsfffsff":
        (router.mark_worker_failure if ev == "f" else router.mark_worker_success)("w0")
        trans.append(router.workers["w0"].health.name)
    out["health_transitions"] = {"events": "ffsfffsff", "states": trans}
    dump("route.json", out)


# ----------------------------------------------------------------------------- kv traces
def run_kv_trace(max_size, default_ttl, ops, policy="lru"):
    """ops: list of [op, *args]. Returns results list + final state (all from the reference)."""
    CLOCK[0] = 1000.0
    kv = ref_kv.KVCache(max_size=max_size, eviction_policy=policy, default_ttl=default_ttl)
    res = []
    for op in ops:
        name = op[0]
        try:
            if name == "advance":
                CLOCK[0] += op[1]
                r = None
            elif name == "set":
                kv.set(op[1], op[2], ttl=op[3])
                r = None
            elif name == "get":
                r = kv.get(op[1], "DFLT")
            elif name == "batch_get":
                r = kv.batch_get(op[1])
            elif name == "batch_set":
                kv.batch_set(dict(op[1]), ttl=op[2])
                r = None
            elif name == "delete":
                r = kv.delete(op[1])
            elif name == "contains":
                r = op[1] in kv
            elif name == "len":
                r = len(kv)
            elif name == "clear":
                kv.clear()
                r = None
            elif name == "stats":
                r = kv.get_stats()
            else:
                raise AssertionError(name)
        except StopIteration:
            r = "!StopIteration"
        except ValueError:  # LFU: min() of an empty _access_count (kvstore.py:95)
            r = "!ValueError"
        res.append(r)
    entries = {k: [e.value, e.created_at, e.last_accessed, e.access_count, e.ttl]
               for k, e in kv.cache.items()}
    final = {"entries": entries, "stats": kv.get_stats(), "cache_order": list(kv.cache)}
    if policy == "lru":
        final["lru"] = list(kv._access_order)
    elif policy == "lfu":
        final["lfu"] = [[k, c] for k, c in kv._access_count.items()]
    out = {"max_size": max_size, "default_ttl": default_ttl, "ops": ops, "results": res, "final": final}
    if policy != "lru":
        out["policy"] = policy
    return out


def random_kv_ops(seed, n_ops, n_keys, with_batch_set, ttl_choices):
    rng = random.Random(seed)
    keys = [f"k{i}" for i in range(n_keys)]
    ops, val, cleared = [], 0, False
    for _ in range(n_ops):
        x = rng.random()
        if x < 0.30:
            val += 1
            ops.append(["set", rng.choice(keys), val, rng.choice(ttl_choices)])
        elif x < 0.55:
            ops.append(["get", rng.choice(keys)])
        elif x < 0.67:
            ops.append(["batch_get", [rng.choice(keys) for _ in range(rng.randrange(0, 12))]])
        elif x < 0.75 and with_batch_set:
            items = []
            for _ in range(rng.randrange(0, 6)):
                val += 1
                items.append([rng.choice(keys), val])
            ops.append(["batch_set", items, rng.choice(ttl_choices)])
        elif x < 0.80:
            ops.append(["delete", rng.choice(keys)])
        elif x < 0.85:
            ops.append(["contains", rng.choice(keys)])
        elif x < 0.88:
            ops.append(["len"])
        elif x < 0.8805 and len(ops) > n_ops // 2 and not cleared:
            cleared = True
            ops.append(["clear"])
        elif x < 0.91:
            ops.append(["stats"])
        else:
            ops.append(["advance", rng.choice([0.0, 0.5, 1.0, 3.0, 10.0])])
    ops.append(["stats"])
    return ops


def gen_kv():
    traces = {}
    # Appendix A.2 cases, by construction
    traces["a2_1_set_A_then_H"] = run_kv_trace(3, None, [["set", "H", 1, None], ["set", "B", 2, None], ["set", "C", 3, None],
                                                        ["set", "A", 4, None], ["set", "H", 5, None], ["stats"]])
    traces["a2_1_set_H_then_A"] = run_kv_trace(3, None, [["set", "H", 1, None], ["set", "B", 2, None], ["set", "C", 3, None],
                                                        ["set", "H", 5, None], ["set", "A", 4, None], ["stats"]])
    traces["a2_2_batch_set_empty"] = run_kv_trace(3, None, [["batch_set", [[f"k{i}", i] for i in range(5)], None], ["stats"]])
    traces["a2_3_batch_set_full"] = run_kv_trace(3, None, [["set", "H", 1, None], ["set", "B", 2, None], ["set", "C", 3, None],
                                                          ["batch_set", [["X", 7], ["Y", 8]], None], ["stats"]])
    traces["a2_3_batch_set_overflow"] = run_kv_trace(3, None, [["set", "H", 1, None], ["set", "B", 2, None], ["set", "C", 3, None],
                                                              ["batch_set", [["X", 7], ["Y", 8], ["Z", 9], ["W", 10]], None],
                                                              ["stats"], ["set", "Q", 11, None], ["get", "X"], ["stats"]])
    traces["a2_3_max_size_0"] = run_kv_trace(0, None, [["set", "a", 1, None], ["stats"]])
    traces["a2_4_ttl_strict_evict_ignores_ttl"] = run_kv_trace(2, None, [
        ["set", "a", 1, 10.0], ["set", "b", 2, None], ["get", "a"], ["advance", 10.0], ["get", "a"],
        ["advance", 1.0], ["set", "c", 3, None], ["get", "a"], ["stats"]])
    traces["a2_5_dups_and_none"] = run_kv_trace(4, None, [["set", "n", None, None], ["get", "n"], ["set", "a", 1, None],
                                                         ["get", "a"], ["get", "a"], ["get", "a"],
                                                         ["batch_get", ["a", "a", "zz"]], ["stats"]])
    traces["a2_6_overwrite"] = run_kv_trace(3, None, [["set", "a", 1, None], ["set", "b", 2, None], ["get", "a"],
                                                     ["advance", 2.0], ["set", "a", 9, None], ["stats"]])
    traces["a2_7_ttl_zero_default"] = run_kv_trace(3, 5.0, [["set", "a", 1, None], ["set", "z", 2, 0], ["get", "z"],
                                                           ["advance", 0.5], ["get", "z"], ["advance", 4.5], ["get", "a"],
                                                           ["advance", 0.5], ["get", "a"], ["stats"]])
    traces["a2_9_contains_len"] = run_kv_trace(4, None, [["set", "a", 1, 1.0], ["set", "b", 2, 1.0], ["set", "c", 3, None],
                                                        ["advance", 2.0], ["contains", "a"], ["stats"], ["len"], ["stats"]])
    traces["a2_10_ref_test_lru"] = run_kv_trace(3, None, [["set", "key1", 1, None], ["set", "key2", 2, None], ["set", "key3", 3, None],
                                                         ["get", "key1"], ["set", "key4", 4, None], ["get", "key2"], ["get", "key1"],
                                                         ["get", "key3"], ["get", "key4"], ["len"]])
    # random differential traces
    cfgs = [(1, 3, None, 600, 8, True, [None, None, 1.0, 5.0]),
            (2, 8, 4.0, 900, 20, True, [None, None, 0, 0.5, 2.0, 20.0]),
            (3, 16, None, 1200, 40, False, [None]),
            (8, 24, 6.0, 1500, 60, True, [None, None, 2.0, 50.0]),
            (9, 100, None, 2500, 260, False, [None, None, 8.0]),
            (4, 5, 2.0, 800, 12, True, [None, 1.0]),
            (5, 64, 30.0, 2000, 200, False, [None, None, None, 3.0]),
            (6, 1, None, 200, 4, True, [None, 1.0]),
            (7, 32, None, 1500, 48, False, [None, 2.0])]
    for seed, ms, dttl, n_ops, n_keys, wbs, ttls in cfgs:
        traces[f"random_{seed}"] = run_kv_trace(ms, dttl, random_kv_ops(seed, n_ops, n_keys, wbs, ttls))
    dump("kv_traces.json", traces)


def gen_kv_policies():
    """LFU and FIFO (src/kvstore.py:67-69,79-80,94-102): constructed cases + random differential traces."""
    traces = {}
    fill3 = [["set", "H", 1, None], ["set", "B", 2, None], ["set", "C", 3, None]]
    for pol in ("lfu", "fifo"):
        traces[f"{pol}_evict_basic"] = run_kv_trace(3, None, fill3 + [["get", "H"], ["get", "H"], ["get", "C"], ["set", "A", 4, None],
                                                                     ["stats"], ["set", "D", 5, None], ["get", "B"], ["get", "A"],
                                                                     ["stats"]], pol)
        traces[f"{pol}_overwrite_moves"] = run_kv_trace(3, None, fill3 + [["get", "B"], ["set", "H", 9, None], ["set", "X", 4, None],
                                                                         ["set", "Y", 5, None], ["stats"]], pol)
        traces[f"{pol}_batch_set_empty"] = run_kv_trace(3, None, [["batch_set", [[f"k{i}", i] for i in range(5)], None], ["stats"]], pol)
        traces[f"{pol}_batch_set_full"] = run_kv_trace(3, None, fill3 + [["get", "H"], ["batch_set", [["X", 7], ["Y", 8]], None],
                                                                        ["stats"], ["batch_set", [["P", 1], ["Q", 2], ["R", 3], ["S", 4]], None],
                                                                        ["stats"], ["get", "S"], ["set", "T", 5, None], ["stats"]], pol)
        traces[f"{pol}_max_size_0"] = run_kv_trace(0, None, [["set", "a", 1, None], ["stats"]], pol)
        traces[f"{pol}_ttl"] = run_kv_trace(2, None, [["set", "a", 1, 10.0], ["set", "b", 2, None], ["get", "a"], ["advance", 11.0],
                                                     ["set", "c", 3, None], ["get", "a"], ["get", "b"], ["len"], ["stats"]], pol)
        cfgs = [(11, 3, None, 500, 8, True, [None, None, 1.0, 5.0]),
                (12, 8, 4.0, 800, 20, True, [None, None, 0, 0.5, 2.0, 20.0]),
                (13, 16, None, 1000, 40, False, [None]),
                (14, 24, 6.0, 1200, 60, True, [None, None, 2.0, 50.0]),
                (15, 64, 30.0, 1500, 200, False, [None, None, None, 3.0]),
                (16, 1, None, 200, 4, True, [None, 1.0])]
        for seed, ms, dttl, n_ops, n_keys, wbs, ttls in cfgs:
            traces[f"{pol}_random_{seed}"] = run_kv_trace(ms, dttl, random_kv_ops(seed, n_ops, n_keys, wbs, ttls), pol)
    dump("kv_policy_traces.json", traces)


# ----------------------------------------------------------------------------- batcher
def gen_batcher():
    out = {}

    async def scenario_concurrent():
        seen = []

        async def cb(model, version, inputs):
            seen.append(list(inputs))
            return [f"r:{x}" for x in inputs]

        b = Batcher(max_batch_size=32, max_latency_ms=50.0, batch_callback=cb)
        await b.start()
        futs = await asyncio.gather(*[b.add_request("n", "1", i) for i in range(100)])
        res = [await f for f in futs]
        stats = await b.get_stats()
        await b.stop()
        return {"batches": seen, "results": res, "stats": stats}

    async def scenario_two_keys():
        seen = []

        async def cb(model, version, inputs):
            seen.append([model, list(inputs)])
            return [[model, x] for x in inputs]

        b = Batcher(max_batch_size=32, max_latency_ms=30.0, batch_callback=cb)
        await b.start()
        futs = []
        for i in range(70):
            futs.append(await b.add_request("n" if i % 2 == 0 else "m", "1", i))
        await b.stop()
        res = [await f for f in futs]
        return {"batches": seen, "results": res}

    async def scenario_errors():
        async def short(model, version, inputs):
            return inputs[:-1]

        b = Batcher(max_batch_size=2, max_latency_ms=50.0, batch_callback=short)
        await b.start()
        f1 = await b.add_request("n", "1", 1)
        f2 = await b.add_request("n", "1", 2)
        errs = []
        for f in (f1, f2):
            try:
                await f
            except Exception as e:  # noqa: BLE001
                errs.append([type(e).__name__, str(e)])
        await b.stop()
        b2 = Batcher(max_batch_size=1, max_latency_ms=50.0, batch_callback=None)
        await b2.start()
        f = await b2.add_request("n", "1", 1)
        try:
            await f
        except Exception as e:  # noqa: BLE001
            errs.append([type(e).__name__, str(e)])
        await b2.stop()
        b3 = Batcher()
        try:
            await b3.add_request("n", "1", 1)
        except Exception as e:  # noqa: BLE001
            errs.append([type(e).__name__, str(e)])
        ctor = []
        for kw in ({"max_batch_size": 0}, {"max_latency_ms": 0}):
            try:
                Batcher(**kw)
            except Exception as e:  # noqa: BLE001
                ctor.append([type(e).__name__, str(e)])

        async def boom(model, version, inputs):
            raise KeyError("boom")

        b4 = Batcher(max_batch_size=3, max_latency_ms=20.0, batch_callback=boom)
        await b4.start()
        fs = [await b4.add_request("n", "1", i) for i in range(3)]
        for f in fs:
            try:
                await f
            except Exception as e:  # noqa: BLE001
                errs.append([type(e).__name__, str(e)])
        await b4.stop()
        return {"errors": errs, "ctor": ctor}

    async def scenario_demo():
        # examples/batcher_demo.py:28-57 : 12 requests, batch 5, mock_batch_inference
        ref_mock.asyncio = types.SimpleNamespace(sleep=_no_sleep)
        sizes = []

        async def cb(model, version, inputs):
            sizes.append(len(inputs))
            return await ref_mock.mock_batch_inference(model, version, inputs, latency_ms=0)

        b = Batcher(max_batch_size=5, max_latency_ms=500.0, batch_callback=cb)
        await b.start()
        futs = [await b.add_request("test-model", "1.0", f"request_{i}") for i in range(12)]
        await b.stop()
        res = [await f for f in futs]
        for r in res:
            r.get("metadata", {}).pop("timestamp", None)
        errs = await ref_mock.mock_batch_inference("mm", "2", list(range(10)), latency_ms=0, error_rate=0.25)
        for r in errs:
            r.get("metadata", {}).pop("timestamp", None)
        return {"batch_sizes": sizes, "results": res, "error_injection": errs}

    out["concurrent_100"] = asyncio.run(scenario_concurrent())
    out["two_keys_70"] = asyncio.run(scenario_two_keys())
    out["errors"] = asyncio.run(scenario_errors())
    out["demo_12"] = asyncio.run(scenario_demo())
    dump("batcher.json", out)


# ----------------------------------------------------------------------------- worker / fake model
class VirtualTimeLoop(asyncio.SelectorEventLoop):
    """Event loop whose clock is virtual: when nothing is ready it jumps to the next timer instead of waiting, so the
    reference Batcher's asyncio.sleep(max_latency) timers (batcher.py:152) fire in order at exact, repeatable times."""

    def __init__(self):
        super().__init__()
        self._vt = 0.0

    def time(self):
        # every look at the clock takes a moment, as on a real clock: timers armed one after the other get strictly
        # increasing deadlines and fire in the order they were armed (equal deadlines would leave it to the heap)
        self._vt += 1e-12
        return self._vt

    def _run_once(self):
        if not self._ready and self._scheduled:
            when = self._scheduled[0]._when
            if when > self._vt:
                self._vt = when
        super()._run_once()


def gen_tick_batcher():
    """The reference Batcher driven by arrival ticks on virtual time: which batches it flushes, in which order and why.
    Tick times never coincide with a timer's deadline (ticks on multiples of 1 ms, latencies of x.5 ms)."""
    import src.batcher as ref_batcher

    def run_case(seed, keys, max_batch, latency_ms, n_ticks, max_per_tick, idle_prob):
        rng = random.Random(seed)
        loop = VirtualTimeLoop()
        asyncio.set_event_loop(loop)
        real_time = ref_batcher.time
        ref_batcher.time = types.SimpleNamespace(time=loop.time)
        ticks, flushed = [], []

        async def scenario():
            async def cb(model, version, inputs):
                flushed.append({"key": f"{model}:{version}", "ids": list(inputs), "at": round(loop.time(), 6)})
                return list(inputs)

            b = Batcher(max_batch_size=max_batch, max_latency_ms=latency_ms, batch_callback=cb)
            await b.start()
            t, rid, futs = 0.0, 0, []
            for _ in range(n_ticks):
                t += rng.choice([1, 1, 2, 3, 7, 20]) / 1000.0
                await asyncio.sleep(t - loop.time())       # timers due before the tick fire here, oldest first
                n0 = len(flushed)
                arr = []
                if rng.random() >= idle_prob:
                    for _i in range(rng.randint(1, max_per_tick)):
                        k = rng.choice(keys)
                        arr.append([k, rid])
                        futs.append(await b.add_request(k.split(":")[0], k.split(":")[1], rid))
                        rid += 1
                ticks.append({"now": round(t, 6), "arrivals": arr, "flushed_before": n0})
            stats_open = await b.get_stats()
            await b.stop()
            res = [await f for f in futs]
            assert res == list(range(rid))
            stats = await b.get_stats()
            return stats_open, stats

        try:
            stats_open, stats = loop.run_until_complete(scenario())
        finally:
            ref_batcher.time = real_time
            loop.close()
            asyncio.set_event_loop(None)
        keep = ("total_batches", "total_requests", "total_batched_requests", "pending_requests")
        return {"keys": keys, "max_batch_size": max_batch, "max_latency_ms": latency_ms, "ticks": ticks, "flushed": flushed,
                "stats_before_stop": {k: stats_open[k] for k in keep}, "stats": {k: stats[k] for k in keep}}

    cases = {
        "one_key_small": run_case(1, ["n:1"], 4, 2.5, 60, 3, 0.3),
        "two_keys": run_case(2, ["n:1", "m:1"], 8, 4.5, 120, 6, 0.2),
        "five_keys_bursts": run_case(3, ["a:1", "a:2", "b:1", "c:1", "d:9"], 16, 10.5, 200, 40, 0.4),
        "mostly_latency": run_case(4, ["n:1", "m:2", "k:3"], 32, 1.5, 150, 4, 0.5),
        "mostly_size": run_case(5, ["n:1", "m:2"], 3, 50.5, 100, 20, 0.0),
    }
    dump("tick_batcher.json", cases)


def gen_load_balancer():
    """The reference LoadBalancer through one scripted session per strategy: picks, pinned requests, failures that take
    a worker out, recovery, unregistering, statistics (last_seen stripped)."""
    import src.load_balancer as ref_lb

    def strip(stats):
        return {w: {k: v for k, v in s.items() if k != "last_seen"} for w, s in stats.items()}

    async def session(strategy):
        random.seed(4242)
        lb = ref_lb.LoadBalancer(strategy=ref_lb.LoadBalancerStrategy(strategy), max_failures=2)
        log = [await lb.get_worker()]                               # no workers yet
        for i in range(4):
            lb.register_worker("w%d" % i, "127.0.0.1:%d" % (9100 + i))
        rng = random.Random(7)
        script = []
        for step in range(60):
            r = rng.random()
            if r < 0.55:
                script.append(["get", None])
            elif r < 0.65:
                script.append(["get", "w%d" % rng.randrange(5)])     # pinned, sometimes unknown
            elif r < 0.9:
                script.append(["stats", "w%d" % rng.randrange(4), rng.random() < 0.6, round(rng.random() / 10, 4)])
            elif r < 0.95:
                script.append(["conn", "w%d" % rng.randrange(4), rng.randrange(5)])
            else:
                script.append(["unregister" if rng.random() < 0.5 else "register", "w%d" % rng.randrange(4)])
        for op in script:
            if op[0] == "get":
                got = await lb.get_worker(op[1])
                log.append(list(got) if got else None)
            elif op[0] == "stats":
                await lb.update_stats(op[1], op[2], op[3])
                log.append("ok")
            elif op[0] == "conn":
                if op[1] in lb.worker_stats:
                    lb.worker_stats[op[1]].active_connections = op[2]
                log.append("ok")
            elif op[0] == "unregister":
                log.append(lb.unregister_worker(op[1]))
            else:
                lb.register_worker(op[1], "127.0.0.1:9999")
                log.append("ok")
        return {"script": script, "log": log, "stats": strip(lb.get_all_stats()), "unknown": lb.get_worker_stats("nope"),
                "repr": repr(lb)}

    out = {s: asyncio.run(session(s)) for s in ("round_robin", "least_connections", "random", "least_latency")}
    dump("load_balancer.json", out)


def gen_worker():
    ref_fake.asyncio = types.SimpleNamespace(sleep=_no_sleep)

    async def run():
        w = Worker("w1")
        w.load_model(ModelConfig(model_name="test-model", model_path="/p", batch_size=8, max_batch_size=32))
        reqs = [{"model": "test-model", "inputs": "hello"}, {"model": "test-model", "inputs": [1, 2, 3]},
                {"model": "test-model", "inputs": (4, 5)}, {"model": "nope", "inputs": 1}, {"inputs": 1},
                {"model": "test-model"}, "notadict", {"model": "test-model", "inputs": {"a": 1}},
                {"model": "test-model", "inputs": 0}, {"model": "", "inputs": 1}]
        outs = []
        for r in reqs:
            o = await w._process_request(r)
            if isinstance(o.get("outputs"), dict):
                o["outputs"]["metadata"].pop("timestamp", None)
                o["outputs"]["metadata"].pop("request_id", None)
            outs.append(o)
        m = w.models["test-model"].get_metrics()
        return {"requests": [list(r) if isinstance(r, tuple) else r for r in
                             [dict(r, inputs=list(r["inputs"])) if isinstance(r, dict) and isinstance(r.get("inputs"), tuple) else r
                              for r in reqs]],
                "tuple_request_index": 2, "responses": outs,
                "request_count": m["request_count"], "error_count": m["error_count"],
                "unload": [w.unload_model("test-model"), w.unload_model("test-model")]}

    dump("worker.json", asyncio.run(run()))


# ----------------------------------------------------------------------------- composed pipeline
def gen_pipeline():
    """Chain the reference's own Router + KVCache + Batcher + FakeModel per SURVEY.md call-stack (E)."""
    ref_fake.asyncio = types.SimpleNamespace(sleep=_no_sleep)
    from src.mock_models import FakeModel

    def run_case(seed, n_shards, max_size, default_ttl, n_batches, batch, id_space, ragged, n_ranks=1,
                 unhealthy=()):
        rng = random.Random(seed)
        reg = make_registry(range(n_shards))
        router = Router(reg)
        for s in range(n_shards):
            router.register_worker(f"w{s}", f"127.0.0.1:{9000 + s}")
            router.mark_worker_success(f"w{s}")
        for s in unhealthy:
            for _ in range(3):
                router.mark_worker_failure(f"w{s}")
        caches = [ref_kv.KVCache(max_size=max_size, eviction_policy="lru", default_ttl=default_ttl)
                  for _ in range(n_shards)]
        fake = FakeModel(ModelConfig(model_name="m", model_path="/p"))
        shards = reg.get_model_version("m", "1").shards

        async def cb(model, version, inputs_list):
            return [(await fake.predict(x))["output"] for x in inputs_list]

        async def infer(inputs_list):
            b = Batcher(max_batch_size=max(1, len(inputs_list)), max_latency_ms=1000.0, batch_callback=cb)
            await b.start()
            futs = [await b.add_request("m", "1", x) for x in inputs_list]
            await b.stop()
            return [await f for f in futs]

        CLOCK[0] = 5000.0
        steps = []
        for _ in range(n_batches):
            CLOCK[0] += rng.choice([0.0, 0.25, 1.0, 4.0])
            per_rank = []
            for _r in range(n_ranks):
                keys = ["user-%010d" % rng.randrange(id_space) for _ in range(batch)]
                if ragged:
                    pays = [[rng.randrange(-1000, 1000) for _ in range(rng.randrange(1, 9))] for _ in range(batch)]
                else:
                    pays = [[rng.randrange(-1000, 1000) for _ in range(4)] for _ in range(batch)]
                per_rank.append([keys, pays])
            flat_keys = [k for ks, _ in per_rank for k in ks]
            flat_pay = [p for _, ps in per_rank for p in ps]
            n = len(flat_keys)
            shard = []
            for k in flat_keys:
                sh = router.route_request("m", "1", k)
                shard.append(-1 if sh is None else shards.index(sh))
            hit, out = [0] * n, [None] * n
            for s in range(n_shards):
                idx = [i for i in range(n) if shard[i] == s]
                miss = []
                for i in idx:
                    v = caches[s].get(flat_keys[i])
                    if v is None:
                        miss.append(i)
                    else:
                        hit[i], out[i] = 1, v
                if miss:
                    res = asyncio.run(infer([flat_pay[i] for i in miss]))
                    for i, r in zip(miss, res):
                        out[i] = r
                    for i in miss:
                        caches[s].set(flat_keys[i], out[i])
            steps.append({"now": CLOCK[0], "per_rank": per_rank, "shard": shard, "hit": hit, "out": out,
                          "state": [{"lru": list(c._access_order), **c.get_stats()} for c in caches]})
        return {"n_shards": n_shards, "max_size": max_size, "default_ttl": default_ttl, "n_ranks": n_ranks,
                "unhealthy": list(unhealthy), "steps": steps}

    cases = {
        "one_shard_small": run_case(1, 1, 8, None, 12, 16, 24, False),
        "one_shard_ttl": run_case(2, 1, 16, 3.0, 16, 24, 40, True),
        "four_shards": run_case(3, 4, 6, 10.0, 10, 32, 64, True),
        "three_shards_failover": run_case(4, 3, 8, None, 8, 24, 48, True, unhealthy=(1,)),
        "two_ranks_two_shards": run_case(5, 2, 10, 5.0, 8, 12, 40, True, n_ranks=2),
        "eight_shards": run_case(6, 8, 4, None, 6, 64, 100, False),
    }
    dump("pipeline.json", cases)


if __name__ == "__main__":
    gen_route()
    gen_kv()
    gen_kv_policies()
    gen_batcher()
    gen_tick_batcher()
    gen_load_balancer()
    gen_worker()
    gen_pipeline()
