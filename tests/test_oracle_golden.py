"""Pins the CPU oracle (oracle/ref_port.py, oracle/pipeline.py) against fixtures produced by the
reference's own code (tests/golden/make_golden.py).  CPU only."""
import asyncio

import pytest

from oracle import ref_port as rp
from oracle.pipeline import BaselinePipeline, OraclePipeline
from conftest import load_golden, replay_ticks


# ----------------------------------------------------------------------------- routing
def test_md5_and_shard_index_match_reference():
    g = load_golden("route.json")
    keys = g["keys"]
    assert [rp.md5_digest(k).hex() for k in keys] == g["md5"]
    for n, want in g["shard_index"].items():
        assert [rp.shard_index(k, int(n)) for k in keys] == want, n


def test_shard_is_list_position_not_shard_id():
    g = load_golden("route.json")
    reg = rp.PortRegistry()
    reg.register_model("m", "1")
    for sid in g["out_of_order"]["shard_ids"]:
        reg.add_shard("m", "1", sid, f"w{sid}")
    got = [reg.shard_for_key("m", "1", k)[0] for k in g["keys"]]
    assert got == g["out_of_order"]["routed_ids"]
    assert reg.shard_for_key("nope", "1", "k") is None and g["unknown_model"]
    reg.register_model("e", "1")
    assert reg.shard_for_key("e", "1", "k") is None and g["no_shards"]
    with pytest.raises(ValueError):
        reg.add_shard("zzz", "1", 0, "w")


def build_port_router(case):
    reg = rp.PortRegistry()
    reg.register_model("m", "1")
    for s in range(case["n"]):
        reg.add_shard("m", "1", s, f"w{s}")
    router = rp.PortRouter(reg, failover_enabled=case["failover_enabled"])
    for s, st in enumerate(case["states"]):
        if st == "UNREGISTERED":
            continue
        router.register_worker(f"w{s}")
        if st == "HEALTHY":
            router.mark_worker_success(f"w{s}")
        elif st == "UNHEALTHY":
            for _ in range(3):
                router.mark_worker_failure(f"w{s}")
    return router


def test_failover_table_matches_reference():
    g = load_golden("route.json")
    for case in g["failover"]:
        router = build_port_router(case)
        got = []
        for k in g["keys"][:48]:
            s = router.route("m", "1", k, case["prefer_healthy"])
            got.append(-1 if s is None else s[0])
        assert got == case["routed"], case["states"]


def test_health_transitions():
    g = load_golden("route.json")["health_transitions"]
    reg = rp.PortRegistry()
    reg.register_model("m", "1")
    reg.add_shard("m", "1", 0, "w0")
    r = rp.PortRouter(reg)
    r.register_worker("w0")
    states = [r.workers["w0"][0]]
    for ev in g["events"]:
        (r.mark_worker_failure if ev == "f" else r.mark_worker_success)("w0")
        states.append(r.workers["w0"][0])
    assert states == g["states"]


# ----------------------------------------------------------------------------- kv
def replay_kv_trace_on_port(t):
    clock = [1000.0]
    kv = rp.PortKVCache(t["max_size"], t["default_ttl"], clock=lambda: clock[0], policy=t.get("policy", "lru"))
    res = []
    for op in t["ops"]:
        name = op[0]
        try:
            if name == "advance":
                clock[0] += op[1]
                r = None
            elif name == "set":
                r = kv.set(op[1], op[2], ttl=op[3])
            elif name == "get":
                r = kv.get(op[1], "DFLT")
            elif name == "batch_get":
                r = kv.batch_get(op[1])
            elif name == "batch_set":
                r = kv.batch_set(dict((k, v) for k, v in op[1]), ttl=op[2])
            elif name == "delete":
                r = kv.delete(op[1])
            elif name == "contains":
                r = kv.contains(op[1])
            elif name == "len":
                r = kv.length()
            elif name == "clear":
                r = kv.clear()
            elif name == "stats":
                r = kv.stats()
        except StopIteration:
            r = "!StopIteration"
        except ValueError:
            r = "!ValueError"
        res.append(r)
    return kv, res


@pytest.mark.parametrize("name", sorted(load_golden("kv_traces.json").keys()))
def test_kv_trace_matches_reference(name):
    t = load_golden("kv_traces.json")[name]
    kv, res = replay_kv_trace_on_port(t)
    assert res == t["results"]
    assert kv.lru_order() == t["final"]["lru"]
    assert kv.stats() == t["final"]["stats"]
    got = {k: [e.value, e.created_at, e.last_accessed, e.access_count, e.ttl] for k, e in kv.cache.items()}
    assert got == t["final"]["entries"]
    assert list(got) == list(t["final"]["entries"])  # dict insertion order too


@pytest.mark.parametrize("name", sorted(load_golden("kv_policy_traces.json").keys()))
def test_kv_policy_trace_matches_reference(name):
    """LFU / FIFO traces recorded from the reference's KVCache (kvstore.py:67-69,79-80,94-102)."""
    t = load_golden("kv_policy_traces.json")[name]
    kv, res = replay_kv_trace_on_port(t)
    assert res == t["results"]
    assert kv.stats() == t["final"]["stats"]
    got = {k: [e.value, e.created_at, e.last_accessed, e.access_count, e.ttl] for k, e in kv.cache.items()}
    assert got == t["final"]["entries"] and list(got) == t["final"]["cache_order"]
    if t["policy"] == "lfu":
        assert [[k, c] for k, c in kv.counts.items()] == t["final"]["lfu"]


# ----------------------------------------------------------------------------- batcher / worker
def test_batcher_scenarios_match_reference():
    g = load_golden("batcher.json")

    async def concurrent():
        seen = []

        async def cb(m, v, inputs):
            seen.append(list(inputs))
            return [f"r:{x}" for x in inputs]

        b = rp.PortBatcher(32, 50.0, cb)
        await b.start()
        futs = await asyncio.gather(*[b.add_request("n", "1", i) for i in range(100)])
        res = [await f for f in futs]
        st = await b.get_stats()
        await b.stop()
        return seen, res, st

    seen, res, st = asyncio.run(concurrent())
    assert seen == g["concurrent_100"]["batches"]
    assert res == g["concurrent_100"]["results"]
    assert st == g["concurrent_100"]["stats"]

    async def two_keys():
        seen = []

        async def cb(m, v, inputs):
            seen.append([m, list(inputs)])
            return [[m, x] for x in inputs]

        b = rp.PortBatcher(32, 30.0, cb)
        await b.start()
        futs = [await b.add_request("n" if i % 2 == 0 else "m", "1", i) for i in range(70)]
        await b.stop()
        return seen, [await f for f in futs]

    seen, res = asyncio.run(two_keys())
    assert seen == g["two_keys_70"]["batches"]
    assert res == g["two_keys_70"]["results"]


TICK_CASES = ["one_key_small", "two_keys", "five_keys_bursts", "mostly_latency", "mostly_size"]


@pytest.mark.parametrize("name", TICK_CASES)
def test_tick_batcher_matches_reference_on_virtual_time(name):
    """PortTickBatcher (the schedule csrc/batcher.cu implements) against the reference's own Batcher run on a
    virtual-time event loop (tests/golden/make_golden.py::gen_tick_batcher)."""
    case = load_golden("tick_batcher.json")[name]
    b = rp.PortTickBatcher(case["max_batch_size"], case["max_latency_ms"])
    state = {}

    def tick(now, arrivals, force):
        if force:
            state["before_stop"] = {"total_batches": b.total_batches, "total_requests": b.total_requests,
                                    "total_batched_requests": b.total_batched_requests,
                                    "pending_requests": sum(b.pending().values())}
        return b.tick(now, arrivals, force)

    replay_ticks(case, tick)
    assert state["before_stop"] == case["stats_before_stop"]
    assert {"total_batches": b.total_batches, "total_requests": b.total_requests,
            "total_batched_requests": b.total_batched_requests, "pending_requests": 0} == case["stats"]


def test_batcher_errors_match_reference():
    g = load_golden("batcher.json")["errors"]

    async def run():
        errs = []

        async def short(m, v, inputs):
            return inputs[:-1]

        b = rp.PortBatcher(2, 50.0, short)
        await b.start()
        fs = [await b.add_request("n", "1", 1), await b.add_request("n", "1", 2)]
        for f in fs:
            try:
                await f
            except Exception as e:  # noqa: BLE001
                errs.append([type(e).__name__, str(e)])
        await b.stop()
        b2 = rp.PortBatcher(1, 50.0, None)
        await b2.start()
        f = await b2.add_request("n", "1", 1)
        try:
            await f
        except Exception as e:  # noqa: BLE001
            errs.append([type(e).__name__, str(e)])
        await b2.stop()
        try:
            await rp.PortBatcher().add_request("n", "1", 1)
        except Exception as e:  # noqa: BLE001
            errs.append([type(e).__name__, str(e)])

        async def boom(m, v, inputs):
            raise KeyError("boom")

        b4 = rp.PortBatcher(3, 20.0, boom)
        await b4.start()
        fs = [await b4.add_request("n", "1", i) for i in range(3)]
        for f in fs:
            try:
                await f
            except Exception as e:  # noqa: BLE001
                errs.append([type(e).__name__, str(e)])
        await b4.stop()
        return errs

    assert asyncio.run(run()) == g["errors"]
    ctor = []
    for kw in ({"max_batch_size": 0}, {"max_latency_ms": 0}):
        try:
            rp.PortBatcher(**kw)
        except Exception as e:  # noqa: BLE001
            ctor.append([type(e).__name__, str(e)])
    assert ctor == g["ctor"]


def test_mock_inference_and_demo_match_reference():
    g = load_golden("batcher.json")["demo_12"]

    async def run():
        sizes = []

        async def cb(m, v, inputs):
            sizes.append(len(inputs))
            return await rp.port_mock_batch_inference(m, v, inputs, latency_ms=0)

        b = rp.PortBatcher(5, 500.0, cb)
        await b.start()
        futs = [await b.add_request("test-model", "1.0", f"request_{i}") for i in range(12)]
        await b.stop()
        res = [await f for f in futs]
        errs = await rp.port_mock_batch_inference("mm", "2", list(range(10)), latency_ms=0, error_rate=0.25)
        return sizes, res, errs

    sizes, res, errs = asyncio.run(run())
    assert sizes == g["batch_sizes"]
    assert res == g["results"]
    assert errs == g["error_injection"]


def test_worker_matches_reference():
    g = load_golden("worker.json")

    async def run():
        w = rp.PortWorker("w1")
        w.load_model("test-model")
        outs = []
        for i, r in enumerate(g["requests"]):
            if i == g["tuple_request_index"]:
                r = dict(r, inputs=tuple(r["inputs"]))
            o = await w.process_request(r)
            if isinstance(o.get("outputs"), dict) and isinstance(o["outputs"]["output"], tuple):
                o["outputs"]["output"] = list(o["outputs"]["output"])
            outs.append(o)
        return w, outs

    w, outs = asyncio.run(run())
    assert outs == g["responses"]
    assert w.models["test-model"].request_count == g["request_count"]
    assert [w.unload_model("test-model"), w.unload_model("test-model")] == g["unload"]


# ----------------------------------------------------------------------------- composed pipeline
@pytest.mark.parametrize("name", sorted(load_golden("pipeline.json").keys()))
@pytest.mark.parametrize("cls", [OraclePipeline, BaselinePipeline])
def test_pipeline_matches_reference_composition(name, cls):
    c = load_golden("pipeline.json")[name]
    p = cls(c["n_shards"], c["max_size"], c["default_ttl"], model=lambda rows: list(rows))
    for s in c["unhealthy"]:
        for _ in range(3):
            p.router.mark_worker_failure(f"w{s}")
    for step in c["steps"]:
        outs = p.process_exchange([(k, pay) for k, pay in step["per_rank"]], step["now"])
        shard = [x for o in outs for x in o["shard"]]
        hit = [x for o in outs for x in o["hit"]]
        out = [x for o in outs for x in o["out"]]
        assert shard == step["shard"]
        assert hit == step["hit"]
        assert out == step["out"]
        assert p.state() == step["state"]
