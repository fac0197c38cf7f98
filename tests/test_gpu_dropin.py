"""GPU: the reference-facing classes used the way the reference's own tests and demos use them."""
import asyncio

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def test_reference_kvstore_unittest_vectors():
    """tests/test_kvstore.py of the reference: test_lru_eviction (:60-79), test_ttl (:45-58, clock injected),
    test_serialization (:118-135), test_clear (:107-116), test_create_kv_store (:165-177)."""
    from die.kvstore import KVStore, create_kv_store
    kv = create_kv_store(max_size=3)
    for i in (1, 2, 3):
        kv.set(f"key{i}", f"value{i}")
    kv.get("key1")
    kv.set("key4", "value4")
    assert kv.get("key2") is None and kv.get("key1") == "value1"
    assert kv.get("key3") == "value3" and kv.get("key4") == "value4" and len(kv) == 3

    t = [100.0]
    kv = KVStore(max_size=3, clock=lambda: t[0])
    kv.set("key1", "value1", ttl=0.1)
    assert kv.get("key1") == "value1"
    t[0] += 0.2
    assert kv.get("key1") is None and "key1" not in kv
    kv.set("key2", "value2", ttl=None)
    assert kv.get("key2") == "value2"

    kv = create_kv_store(max_size=3)
    for i, v in enumerate(["string", 123, 3.14, True, None, {"key": "value"}, [1, 2, 3], (4, 5, 6), {"a", "b"}]):
        kv.set(f"key{i}", v)
        assert kv.get(f"key{i}") == v
    kv.clear()
    assert len(kv) == 0 and kv.get("key1") is None
    assert create_kv_store().max_size == 1000 and isinstance(create_kv_store(max_size=500), KVStore)
    assert kv.get("nonexistent", "default") == "default" and kv.delete("nonexistent") is False


def test_batcher_device_path_packs_runs_and_unpacks_in_arrival_order():
    from die import Batcher, FakeModel, ModelConfig, PackedBatch
    model = FakeModel(ModelConfig("echo", "/p"), simulate_latency=False)
    seen = []

    async def device_cb(name, version, packed: PackedBatch):
        seen.append((tuple(packed.data.shape), packed.lengths.tolist()))
        return model.forward(packed)

    async def run():
        rng = np.random.default_rng(4)
        b = Batcher(max_batch_size=32, max_latency_ms=30.0, device_callback=device_cb)
        await b.start()
        payloads = [rng.standard_normal(int(n)).astype(np.float32) for n in rng.integers(1, 50, 70)]
        futs = [await b.add_request("echo", "1", p) for p in payloads]
        await b.stop()
        return payloads, [await f for f in futs]

    payloads, results = asyncio.run(run())
    assert [s[0][0] for s in seen] == [32, 32, 6]
    for p, r in zip(payloads, results):
        assert np.array_equal(r.cpu().numpy(), p)
    lens = [len(p) for p in payloads]
    assert seen[0][1] == lens[:32] and seen[2][1] == lens[64:]


def test_batcher_device_path_error_fans_out():
    from die import Batcher

    async def bad(name, version, packed):
        raise RuntimeError("device callback failed")

    async def run():
        b = Batcher(2, 20.0, device_callback=bad)
        await b.start()
        fs = [await b.add_request("m", "1", [1.0, 2.0]), await b.add_request("m", "1", [3.0])]
        errs = []
        for f in fs:
            try:
                await f
            except RuntimeError as e:
                errs.append(str(e))
        await b.stop()
        return errs

    assert asyncio.run(run()) == ["device callback failed"] * 2


def test_worker_and_models_on_device_tensors():
    from die import LinearModel, ModelConfig, Worker
    from die.mock_models import FakeModel
    from oracle import models
    cfg = ModelConfig("lin", "/p")
    rng = np.random.default_rng(0)
    w = (rng.standard_normal((128, 128)) / 11.3).astype(np.float32)
    b = rng.standard_normal(128).astype(np.float32)
    wk = Worker("w0")
    assert wk.load_model(cfg, LinearModel(cfg, torch.from_numpy(w).cuda(), torch.from_numpy(b).cuda(), simulate_latency=False))
    x = rng.standard_normal((64, 128)).astype(np.float32)
    resp = asyncio.run(wk._process_request({"model": "lin", "inputs": torch.from_numpy(x).cuda()}))
    assert resp["success"] and resp["worker_id"] == "w0" and resp["outputs"]["metadata"]["batch_size"] == 1
    y = resp["outputs"]["output"].cpu().numpy()
    assert np.all(np.abs(y - models.linear_f64(x, w, b)) <= 1e-5 * models.linear_abs_bound(x, w) + 1e-6)
    echo = FakeModel(ModelConfig("e", "/p"), simulate_latency=False)
    t = torch.randn(1000, device="cuda")
    out = asyncio.run(echo.predict(t))
    assert torch.equal(out["output"], t) and out["model"] == "e"
    assert asyncio.run(wk._process_request({"model": "nope", "inputs": 1})) == {"error": "Model 'nope' not found", "success": False}
