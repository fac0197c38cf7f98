"""CPU tier: the C-ABI library loads and exports what include/die.h declares, the product's MD5
code (host build) and the host-side logic of the drop-in classes agree with the reference goldens.
No compute entry point is called here (there is no GPU)."""
import asyncio
import hashlib
import os
import random
import re
import sys

import numpy as np
import pytest

from conftest import PKG_DIR, ROOT, load_golden
from oracle import c_oracle
from oracle import ref_port as rp


def test_library_exports_every_declared_symbol():
    from die import _cabi
    lib = _cabi.load()
    header = open(os.path.join(ROOT, "include", "die.h")).read()
    declared = set(re.findall(r"^(?:die_status_t|int32_t|uint64_t|const char\*|die_kv_t\*)\s+(die_[a-z0-9_]+)\s*\(",
                              header, re.M))
    assert declared, "no declarations parsed"
    assert declared == set(_cabi.SIGNATURES), declared ^ set(_cabi.SIGNATURES)
    for name in declared:
        assert hasattr(lib, name), name
    assert lib.die_abi_version() == 1
    assert lib.die_launch_count() == 0


def test_compute_without_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from die import KVCache, ModelRegistry
    with pytest.raises(RuntimeError, match="no CUDA device"):
        KVCache(max_size=4)
    reg = ModelRegistry()
    reg.register_model("m", "1", "/p", {}, {})
    reg.add_shard("m", "1", 0, "w0")
    with pytest.raises(RuntimeError, match="no CUDA device"):
        reg.get_shard_for_key("m", "1", "k")


def test_product_never_imports_the_oracle():
    for dirpath, _, files in os.walk(PKG_DIR):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                text = open(os.path.join(dirpath, f), errors="replace").read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, re.M), f
                for needle in ("liboracle", "c_oracle", "ref_port", "libmd5kernelhost", "import_module(\"oracle"):
                    assert needle not in text, (f, needle)


def test_kernel_md5_host_build_matches_hashlib():
    rng = random.Random(3)
    keys = ["".join(chr(rng.randrange(32, 127)) for _ in range(n)) for n in range(0, 300)]
    keys += ["é" * n for n in (1, 27, 28, 32, 64)] + ["user-%010d" % rng.randrange(10 ** 7) for _ in range(500)]
    kb, off = c_oracle.encode(keys)
    for n in (1, 2, 3, 8, 1000, 2 ** 31 - 1):
        dig, shard = c_oracle.kernel_md5_host(kb, off, n)
        want = [hashlib.md5(k.encode()).digest() for k in keys]
        assert [d.tobytes() for d in dig] == want
        assert shard.tolist() == [int.from_bytes(w, "big") % n for w in want]


def test_c_oracle_matches_python_port():
    rng = random.Random(5)
    keys = ["k%d" % rng.randrange(10 ** 9) for _ in range(2000)] + ["", "x" * 1000]
    kb, off = c_oracle.encode(keys)
    for n in (1, 3, 8, 97):
        shard, dig = c_oracle.route(kb, off, n)
        assert shard.tolist() == [rp.shard_index(k, n) for k in keys]
        assert [d.tobytes() for d in dig] == [rp.md5_digest(k) for k in keys]


def test_router_health_masks_reproduce_failover_table():
    """Router.health_masks is the host-side lowering the kernel consumes; combined with the C oracle
    of the kernel's decision rule it must reproduce the reference's routing for every golden state."""
    from die import ModelRegistry, Router
    g = load_golden("route.json")
    keys = g["keys"][:48]
    kb, off = c_oracle.encode(keys)
    for case in g["failover"]:
        reg = ModelRegistry()
        reg.register_model("m", "1", "/p", {}, {})
        for s in range(case["n"]):
            reg.add_shard("m", "1", s, f"w{s}")
        router = Router(reg, failover_enabled=case["failover_enabled"])
        for s, st in enumerate(case["states"]):
            if st == "UNREGISTERED":
                continue
            router.register_worker(f"w{s}", f"127.0.0.1:{9000 + s}")
            if st == "HEALTHY":
                router.mark_worker_success(f"w{s}")
            elif st == "UNHEALTHY":
                for _ in range(3):
                    router.mark_worker_failure(f"w{s}")
        routable, healthy = router.health_masks(reg.get_model_version("m", "1").shards, case["prefer_healthy"])
        shard, _ = c_oracle.route(kb, off, case["n"], routable, healthy, case["failover_enabled"])
        assert shard.tolist() == case["routed"], case


def test_router_health_transitions_and_stats():
    from die import ModelRegistry, Router
    g = load_golden("route.json")["health_transitions"]
    router = Router(ModelRegistry())
    router.register_worker("w0", "127.0.0.1:1")
    states = [router.workers["w0"].health.name]
    for ev in g["events"]:
        (router.mark_worker_failure if ev == "f" else router.mark_worker_success)("w0")
        states.append(router.workers["w0"].health.name)
    assert states == g["states"]
    st = router.get_stats()
    assert st["total_workers"] == 1 and st["failover_enabled"] is True
    assert router.get_worker_info("w0")["health"] == states[-1]
    assert router.get_worker_info("nope") is None


def test_registry_metadata_roundtrip():
    from die import ModelRegistry
    reg = ModelRegistry()
    reg.register_model("m", "1", "/p", {"i": "f32"}, {"o": "f32"}, metadata={"d": 1})
    s0 = reg.add_shard("m", "1", 0, "w0", metadata={"gpu": "B200"})
    assert reg.add_shard("m", "1", 0, "other") is s0  # idempotent on shard_id
    reg.add_shard("m", "1", 1, "w1")
    with pytest.raises(ValueError):
        reg.add_shard("zz", "1", 0, "w")
    reg2 = ModelRegistry.from_dict(reg.to_dict())
    assert reg2.to_dict() == reg.to_dict()
    assert sorted(reg2.get_worker_models("w0")) == [("m", "1")]
    assert reg.get_model_hash("m", "1") and reg.get_model_hash("m", "2") is None
    assert reg.list_models() == ["m"] and reg.list_versions("m") == ["1"]
    assert reg.get_shard_for_key("nope", "1", "k") is None  # unknown model: no kernel involved
    reg.register_model("e", "1", "/p", {}, {})
    assert reg.get_shard_for_key("e", "1", "k") is None  # no shards


def test_batcher_host_logic_matches_reference_goldens():
    from die import Batcher
    g = load_golden("batcher.json")

    async def concurrent():
        seen = []

        async def cb(m, v, inputs):
            seen.append(list(inputs))
            return [f"r:{x}" for x in inputs]

        b = Batcher(32, 50.0, cb)
        await b.start()
        futs = await asyncio.gather(*[b.add_request("n", "1", i) for i in range(100)])
        res = [await f for f in futs]
        st = await b.get_stats()
        await b.stop()
        return seen, res, st

    seen, res, st = asyncio.run(concurrent())
    assert seen == g["concurrent_100"]["batches"] and res == g["concurrent_100"]["results"]
    assert st == g["concurrent_100"]["stats"]

    async def two_keys():
        seen = []

        async def cb(m, v, inputs):
            seen.append([m, list(inputs)])
            return [[m, x] for x in inputs]

        b = Batcher(32, 30.0, cb)
        await b.start()
        futs = [await b.add_request("n" if i % 2 == 0 else "m", "1", i) for i in range(70)]
        await b.stop()
        return seen, [await f for f in futs]

    seen, res = asyncio.run(two_keys())
    assert seen == g["two_keys_70"]["batches"] and res == g["two_keys_70"]["results"]


def test_batcher_error_contract_matches_reference():
    from die import Batcher
    g = load_golden("batcher.json")["errors"]

    async def run():
        errs = []

        async def short(m, v, inputs):
            return inputs[:-1]

        b = Batcher(2, 50.0, short)
        await b.start()
        for f in [await b.add_request("n", "1", 1), await b.add_request("n", "1", 2)]:
            try:
                await f
            except Exception as e:  # noqa: BLE001
                errs.append([type(e).__name__, str(e)])
        await b.stop()
        b2 = Batcher(1, 50.0, None)
        await b2.start()
        try:
            await (await b2.add_request("n", "1", 1))
        except Exception as e:  # noqa: BLE001
            errs.append([type(e).__name__, str(e)])
        await b2.stop()
        try:
            await Batcher().add_request("n", "1", 1)
        except Exception as e:  # noqa: BLE001
            errs.append([type(e).__name__, str(e)])

        async def boom(m, v, inputs):
            raise KeyError("boom")

        b4 = Batcher(3, 20.0, boom)
        await b4.start()
        for f in [await b4.add_request("n", "1", i) for i in range(3)]:
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
            Batcher(**kw)
        except Exception as e:  # noqa: BLE001
            ctor.append([type(e).__name__, str(e)])
    assert ctor == g["ctor"]


def test_mock_inference_worker_and_fake_model_match_reference():
    from die import Batcher, ModelConfig, Worker, mock_batch_inference
    from die.mock_models import FakeModel
    g = load_golden("batcher.json")["demo_12"]

    async def demo():
        sizes = []

        async def cb(m, v, inputs):
            sizes.append(len(inputs))
            return await mock_batch_inference(m, v, inputs, latency_ms=0)

        b = Batcher(5, 500.0, cb)
        await b.start()
        futs = [await b.add_request("test-model", "1.0", f"request_{i}") for i in range(12)]
        await b.stop()
        res = [await f for f in futs]
        errs = await mock_batch_inference("mm", "2", list(range(10)), latency_ms=0, error_rate=0.25)
        for r in res + errs:
            r.get("metadata", {}).pop("timestamp", None)
        return sizes, res, errs

    sizes, res, errs = asyncio.run(demo())
    assert sizes == g["batch_sizes"] and res == g["results"] and errs == g["error_injection"]

    w = load_golden("worker.json")

    async def worker():
        wk = Worker("w1")
        cfg = ModelConfig(model_name="test-model", model_path="/p", batch_size=8, max_batch_size=32)
        assert wk.load_model(cfg, FakeModel(cfg, simulate_latency=False)) and wk.load_model(cfg)
        outs = []
        for i, r in enumerate(w["requests"]):
            if i == w["tuple_request_index"]:
                r = dict(r, inputs=tuple(r["inputs"]))
            o = await wk._process_request(r)
            if isinstance(o.get("outputs"), dict):
                o["outputs"]["metadata"].pop("timestamp")
                o["outputs"]["metadata"].pop("request_id")
                if isinstance(o["outputs"]["output"], tuple):
                    o["outputs"]["output"] = list(o["outputs"]["output"])
            outs.append(o)
        return wk, outs

    wk, outs = asyncio.run(worker())
    assert outs == w["responses"]
    assert wk.models["test-model"].request_count == w["request_count"]
    assert [wk.unload_model("test-model"), wk.unload_model("test-model")] == w["unload"]


def test_kv_constructor_surface_without_gpu():
    from die.kvstore import CacheEntry, KVCache, KVStore, create_kv_store
    assert KVStore is KVCache and create_kv_store is KVCache
    e = CacheEntry(value=1, created_at=2.0, last_accessed=3.0)
    assert (e.access_count, e.ttl) == (0, None)


def test_header_is_c99_and_the_c_example_links(tmp_path):
    """The boundary is a C ABI: include/die.h must compile as plain C, and a plain-C caller (examples/pipeline_host.c)
    must build and link against libdie.so.  Without a GPU it has to fail loudly at its first device call."""
    import subprocess
    import torch
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    libdir = os.path.join(root, "distributed-inference-engine_b200", "die", "_lib")
    if not os.path.exists(os.path.join(libdir, "libdie.so")):
        import __graft_entry__ as g
        g.build()
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-pedantic", "-fsyntax-only", "-x", "c",
                    os.path.join(root, "include", "die.h")], check=True)
    exe = str(tmp_path / "pipeline_host")
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Wextra", "-Werror", "-I" + os.path.join(root, "include"),
                    os.path.join(root, "examples", "pipeline_host.c"), "-o", exe, "-L" + libdir, "-ldie",
                    "-Wl,-rpath," + libdir], check=True)
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    if torch.cuda.is_available():
        assert r.returncode == 0, r.stderr
        assert "pass 0 hits: 0 0 0 0 0 0 0 0" in r.stdout and "pass 1 hits: 1 1 1 1 1 1 1 1" in r.stdout
    else:
        assert r.returncode == 1 and "die_set_device" in r.stderr


def test_bench_reference_arm_prints_the_contract_line():
    """`bench.py --impl reference` (the reference's pure-Python path on the host cores) needs no GPU and must print
    one JSON line with the keys the driver reads; under torchrun only rank 0 prints."""
    import json
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1"],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-500:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    for k in ("metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
              "vs_baseline", "dtype", "data", "config", "impl", "cpu_baseline", "e2e"):
        assert k in line, k
    from oracle import refrun
    assert line["impl"] == "reference" and line["value"] > 0
    # the reference's own classes when oracle/_ref holds the verbatim copy (oracle/fetch_ref.py), else the port
    assert line["cpu_baseline"]["kind"] == ("reference" if refrun.available() else "port")
    assert line["config"]["kv_max_size"] == 1 << 20 and line["config"]["id_space"] == 10 ** 7  # the GPU arm's configuration
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and "workload" in line["config"]
    env = dict(os.environ, RANK="1", WORLD_SIZE="2")
    r1 = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--gpus", "2"],
                        capture_output=True, text=True, timeout=120, env=env)
    assert r1.returncode == 0 and r1.stdout.strip() == ""


def test_bench_gpu_arm_has_no_cpu_fallback():
    """Without a GPU the product arm of bench.py must stop with an error, not produce a number some other way."""
    import subprocess
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--steps", "1", "--warmup", "1"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode != 0
    assert not any(line.startswith("{") for line in r.stdout.splitlines())


def test_reference_classes_composed_by_refrun_agree_with_the_port():
    """oracle/refrun.py chains the reference's OWN Router / KVCache / Batcher (verbatim copy under oracle/_ref, made
    by oracle/fetch_ref.py wherever /root/reference exists); the restatement in oracle/pipeline.py must give the
    same shard ids, hit flags, results and per-owner LRU state on a seeded run with evictions and TTL expiry."""
    from oracle import refrun
    from oracle.pipeline import OraclePipeline
    if not refrun.available():
        pytest.skip("oracle/_ref is absent on this box (the reference checkout does not travel)")
    import random
    rng = random.Random(3)
    model = lambda rows: [[2 * v for v in r] for r in rows]  # noqa: E731
    a = refrun.ReferencePipeline(3, 40, 5.0, model, max_batch_size=64)
    b = OraclePipeline(3, 40, 5.0, model)
    now = 10.0
    for _ in range(30):
        now += rng.choice([0.0, 0.5, 3.0, 6.0])
        keys = ["user-%010d" % rng.randrange(200) for _ in range(64)]
        pays = [[rng.randrange(100) for _ in range(3)] for _ in range(64)]
        ra, rb = a.process(keys, pays, now), b.process(keys, pays, now)
        assert ra == rb
    sa, sb = a.state(), b.state()
    for x, y in zip(sa, sb):
        assert x["lru"] == y["lru"]
        assert (x["hits"], x["misses"], x["evictions"], x["size"]) == (y["hits"], y["misses"], y["evictions"], y["size"])
    assert sum(x["evictions"] for x in sa) > 0 and sum(x["hits"] for x in sa) > 0
