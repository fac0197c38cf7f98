"""GPU parity: the composed path (route -> get -> pack -> model -> unpack -> set) through the C-ABI vs
the reference composition goldens and the CPU oracle.  Shard ids, hit flags, LRU state: bit-exact;
echo outputs bit-exact; linear outputs within the fp32 tolerance of oracle/models.py."""
import hashlib

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import models
from oracle.pipeline import OraclePipeline

pytestmark = pytest.mark.gpu
WIDTH = 8


def pad(rows):
    out = np.zeros((len(rows), WIDTH), dtype=np.float32)
    for i, r in enumerate(rows):
        out[i, :len(r)] = r
    return out


def lru_digests(pipe):
    d = pipe.kv.dump()
    return [bytes(x) for x in d["digests"].numpy().view(np.uint8).reshape(-1, 16)]


@pytest.mark.parametrize("name", sorted(load_golden("pipeline.json").keys()))
def test_reference_composition_goldens(name):
    """One DispatchPipeline per shard on this GPU; grouping by shard is done by the test (the NCCL
    exchange is covered elsewhere), owners see sources in rank order as SURVEY.md section 8(e) defines."""
    from die import ModelRegistry, Router, ops
    from die.pipeline import DispatchPipeline
    c = load_golden("pipeline.json")[name]
    n = c["n_shards"]
    reg = ModelRegistry()
    reg.register_model("m", "1", "/p", {}, {})
    router = Router(reg)
    for s in range(n):
        reg.add_shard("m", "1", s, f"w{s}")
        router.register_worker(f"w{s}", f"127.0.0.1:{9000 + s}")
        router.mark_worker_success(f"w{s}")
    for s in c["unhealthy"]:
        for _ in range(3):
            router.mark_worker_failure(f"w{s}")
    pipes = [DispatchPipeline(c["max_size"], 512, WIDTH, model="echo", default_ttl=c["default_ttl"]) for _ in range(n)]
    for step in c["steps"]:
        keys = [k for ks, _ in step["per_rank"] for k in ks]
        pays = [p for _, ps in step["per_rank"] for p in ps]
        kb = ops.KeyBatch.from_strings(keys)
        shard = router.route("m", "1", kb)
        _, dig, _ = ops.route_md5(kb, 1)
        assert shard.tolist() == step["shard"]
        x = torch.from_numpy(pad(pays)).cuda()
        hit_all = np.zeros(len(keys), dtype=np.int64)
        out_all = np.zeros((len(keys), WIDTH), dtype=np.float32)
        sh = shard.cpu().numpy()
        for s in range(n):
            idx = np.nonzero(sh == s)[0]
            if idx.size == 0:
                continue
            it = torch.from_numpy(idx).cuda()
            out, hit = pipes[s].run_owner_device(dig[it].contiguous(), x[it].contiguous(), step["now"])
            hit_all[idx] = hit.cpu().numpy()
            out_all[idx] = out.cpu().numpy()
        assert hit_all.tolist() == step["hit"]
        for i, want in enumerate(step["out"]):
            assert out_all[i, :len(want)].tolist() == [float(v) for v in want], i
        for s in range(n):
            st, want = pipes[s].kv.stats(), step["state"][s]
            assert (st["hits"], st["misses"], st["evictions"], st["size"]) == (want["hits"], want["misses"], want["evictions"], want["size"])
            assert lru_digests(pipes[s]) == [hashlib.md5(k.encode()).digest() for k in want["lru"]]


@pytest.mark.parametrize("name", sorted(load_golden("pipeline.json").keys()))
def test_coordinator_matches_reference_compositions(name):
    """die.Coordinator = the composition the reference's README describes (Router -> per-shard cache -> batcher ->
    model) as product code: same golden runs of the reference's own classes, grouping by shard done by the
    bucket kernel instead of by the test, health-aware routing and failover included."""
    from die import Coordinator, ModelRegistry, Router
    c = load_golden("pipeline.json")[name]
    n = c["n_shards"]
    reg = ModelRegistry()
    reg.register_model("m", "1", "/p", {}, {})
    router = Router(reg)
    for s in range(n):
        reg.add_shard("m", "1", s, f"w{s}")
        router.register_worker(f"w{s}", f"127.0.0.1:{9000 + s}")
        router.mark_worker_success(f"w{s}")
    for s in c["unhealthy"]:
        for _ in range(3):
            router.mark_worker_failure(f"w{s}")
    co = Coordinator(router, "m", "1", c["max_size"], 512, WIDTH, model="echo", default_ttl=c["default_ttl"])
    for step in c["steps"]:
        keys = [k for ks, _ in step["per_rank"] for k in ks]
        pays = [p for _, ps in step["per_rank"] for p in ps]
        out, hit, shard = co.process(keys, torch.from_numpy(pad(pays)).cuda(), step["now"])
        assert shard.tolist() == step["shard"] and hit.tolist() == step["hit"]
        out = out.cpu().numpy()
        for i, want in enumerate(step["out"]):
            assert out[i, :len(want)].tolist() == [float(v) for v in want], i
        for s in range(n):
            st, want = co.pipes[s].kv.stats(), step["state"][s]
            assert (st["hits"], st["misses"], st["evictions"], st["size"]) == (want["hits"], want["misses"], want["evictions"], want["size"])
            assert lru_digests(co.pipes[s]) == [hashlib.md5(k.encode()).digest() for k in want["lru"]]


@pytest.mark.parametrize("model,B,kv_size,space,steps", [("echo", 4096, 20000, 30000, 16), ("linear", 4096, 6000, 9000, 12),
                                                         ("linear", 1000, 300, 500, 20), ("linear_tf32", 4096, 6000, 9000, 10)])
def test_seeded_batches_vs_oracle(model, B, kv_size, space, steps):
    from die.pipeline import DispatchPipeline
    D = 128
    rng = np.random.default_rng(B + kv_size)
    w = (rng.standard_normal((D, D)) / np.sqrt(D)).astype(np.float32)
    b = rng.standard_normal(D).astype(np.float32)
    if model == "echo":
        pipe = DispatchPipeline(kv_size, B, D, model="echo", default_ttl=9.0)
        fn = lambda rows: list(rows)  # noqa: E731
    else:
        pipe = DispatchPipeline(kv_size, B, D, D, model=model, weight=w, bias=b, default_ttl=9.0)

        def fn(rows):
            # value = (result row, its rounding-error scale sum_k |x_k w_k|): a cached result keeps the bound of
            # the payload it was computed from, so hits are held to the stated per-element tolerance as well
            if not rows:
                return []
            xs = np.stack(rows)
            y, bound = models.linear_f64(xs, w, b), models.linear_abs_bound(xs, w)
            return [np.stack([y[i], bound[i]]) for i in range(len(rows))]
    ref = OraclePipeline(1, kv_size, 9.0, fn)
    now = 50.0
    for step in range(steps):
        now += float(rng.choice([0.5, 1.0, 4.0]))
        keys = ["user-%010d" % i for i in rng.integers(0, space, B)]
        x = rng.standard_normal((B, D)).astype(np.float32)
        out, hit = pipe.process(keys, x, now)
        want = ref.process(keys, [x[i] for i in range(B)], now)
        assert hit.tolist() == want["hit"], step
        want_out = np.stack(want["out"])
        if model == "echo":
            assert np.array_equal(out, want_out.astype(np.float32))
            continue
        want_y, bound = want_out[:, 0], want_out[:, 1]
        # oracle/models.py: fp32 FMA path |d| <= 1e-5 * sum|x w| + 1e-6, tf32 inputs |d| <= 2**-10 * sum|x w| + 1e-6
        rel = 1e-5 if model == "linear" else 2.0 ** -10
        assert np.all(np.abs(out.astype(np.float64) - want_y) <= rel * bound + 1e-6), step
    st, rs = pipe.kv.stats(), ref.state()[0]
    assert (st["hits"], st["misses"], st["evictions"], st["size"]) == (rs["hits"], rs["misses"], rs["evictions"], rs["size"])
    assert lru_digests(pipe) == [hashlib.md5(k.encode()).digest() for k in rs["lru"]]
    assert pipe.kv.sync_status()[2] == 0 or kv_size < B  # large tables never need the sequential replay


@pytest.mark.parametrize("B,D,H,kv_size,space,steps", [(700, 256, 512, 2000, 3000, 10), (1024, 512, 256, 600, 1500, 6)])
def test_mlp_bf16_is_a_pipeline_model(B, D, H, kv_size, space, steps):
    """BASELINE.json configs[4] as a path (the model forward of fake_model.py:45-58 replaced by the bf16 tcgen05 MLP):
    route -> probe -> pack the misses -> bf16 MLP -> unpack -> fill, through die_pipeline_run_host and (second table
    smaller than a batch) the general sequence.  Hit flags, counters and LRU order exact; results within
    oracle/models.py's MLP tolerance |d| <= 2**-7 |y| + 1e-2 max_j |y_j| per row, cached rows included."""
    from die.pipeline import DispatchPipeline
    rng = np.random.default_rng(B + D)
    w1 = models.to_bf16((rng.standard_normal((H, D)) / np.sqrt(D)).astype(np.float32))
    w2 = models.to_bf16((rng.standard_normal((D, H)) / np.sqrt(H)).astype(np.float32))
    b1, b2 = rng.standard_normal(H).astype(np.float32), rng.standard_normal(D).astype(np.float32)
    pipe = DispatchPipeline(kv_size, B, D, D, model="mlp_bf16", weight=(w1, w2), bias=(b1, b2), default_ttl=9.0)

    def fn(rows):
        if not rows:
            return []
        y = models.mlp_bf16(models.to_bf16(np.stack(rows)), w1, b1, w2, b2)
        return [y[i] for i in range(len(rows))]
    ref = OraclePipeline(1, kv_size, 9.0, fn)
    now = 50.0
    for step in range(steps):
        now += float(rng.choice([0.5, 1.0, 4.0]))
        keys = ["user-%010d" % i for i in rng.integers(0, space, B)]
        x = rng.standard_normal((B, D)).astype(np.float32)
        out, hit = pipe.process(keys, x, now)
        want = ref.process(keys, [x[i] for i in range(B)], now)
        assert hit.tolist() == want["hit"], step
        y = np.stack(want["out"])
        tol = 2.0 ** -7 * np.abs(y) + 1e-2 * np.abs(y).max(axis=1, keepdims=True)
        assert np.all(np.abs(out.astype(np.float64) - y) <= tol), step
    st, rs = pipe.kv.stats(), ref.state()[0]
    assert (st["hits"], st["misses"], st["evictions"], st["size"]) == (rs["hits"], rs["misses"], rs["evictions"], rs["size"])
    assert st["hits"] > 0 and st["evictions"] > 0
    assert lru_digests(pipe) == [hashlib.md5(k.encode()).digest() for k in rs["lru"]]


def test_device_resident_entry_point_equals_host_entry_point():
    from die import ops
    from die.pipeline import DispatchPipeline
    rng = np.random.default_rng(9)
    B, D = 2048, 64
    a = DispatchPipeline(5000, B, D, model="echo")
    b = DispatchPipeline(5000, B, D, model="echo")
    for step in range(4):
        keys = ["k%d" % i for i in rng.integers(0, 7000, B)]
        x = rng.standard_normal((B, D)).astype(np.float32)
        out_h, hit_h = a.process(keys, x, 1.0 + step)
        out_d, hit_d = b.run_device(ops.KeyBatch.from_strings(keys), torch.from_numpy(x).cuda(), 1.0 + step)
        assert np.array_equal(out_h, out_d.cpu().numpy()) and np.array_equal(hit_h, hit_d.cpu().numpy())


def test_queue_entry_points_equal_per_batch_calls():
    """die_pipeline_run_device_many / run_host_many = the same batches one call at a time."""
    from die import ops
    from die.pipeline import DispatchPipeline
    rng = np.random.default_rng(21)
    n, B, D = 7, 512, 32
    ids = rng.integers(0, 3000, (n, B))
    keys = [["q%08d" % i for i in row] for row in ids]
    x = rng.standard_normal((n, B, D)).astype(np.float32)
    kb = np.stack([np.frombuffer("".join(k).encode(), dtype=np.uint8) for k in keys])
    off = np.arange(B + 1, dtype=np.int64) * 9
    a = DispatchPipeline(1500, B, D, model="echo", default_ttl=3.0)
    b = DispatchPipeline(1500, B, D, model="echo", default_ttl=3.0)
    c = DispatchPipeline(1500, B, D, model="echo", default_ttl=3.0)
    want_out, want_hit = [], []
    for j in range(n):
        o, h = a.process(keys[j], x[j], 5.0 + j * 1.5)
        want_out.append(o)
        want_hit.append(h)
    out = torch.empty((n, B, D), device="cuda")
    hit = torch.empty((n, B), dtype=torch.uint8, device="cuda")
    b.run_device_many(torch.from_numpy(kb).cuda(), torch.from_numpy(off).cuda(), torch.from_numpy(x).cuda(), 5.0, 1.5, out, hit)
    assert np.array_equal(out.cpu().numpy(), np.stack(want_out)) and np.array_equal(hit.cpu().numpy(), np.stack(want_hit))
    hout = torch.empty((n, B, D)).pin_memory()
    hhit = torch.empty((n, B), dtype=torch.uint8).pin_memory()
    kbp = torch.from_numpy(np.concatenate([kb, np.zeros((n, 16), np.uint8)], axis=1)).pin_memory()
    c.run_host_many(kbp, torch.from_numpy(off).pin_memory(), torch.from_numpy(x).pin_memory(), 5.0, 1.5, hout, hhit)
    assert np.array_equal(hout.numpy(), np.stack(want_out)) and np.array_equal(hhit.numpy(), np.stack(want_hit))
    assert a.kv.stats() == b.kv.stats() == c.kv.stats()


@pytest.mark.parametrize("model", ["echo", "linear_tf32"])
def test_graph_replayed_queue_equals_stream_launched_steps(model):
    """A long queue goes through captured replays of 8 steps (csrc/pipeline.cu replay_fast) whose per-step
    arguments come from a device-resident descriptor; a twin pipeline gets the same batches one stream-launched
    call at a time.  Results, counters and the whole table (LRU order, timestamps) must be identical, through
    evictions, TTL expiry and table maintenance."""
    from die import ops
    from die.pipeline import DispatchPipeline
    rng = np.random.default_rng(33)
    n, B, D, max_size = 141, 256, 32, 20000  # 8 replays of 16 + 13 stream steps; the table fills after ~100 batches
    ids = rng.integers(0, 60000, (n, B))
    ids[:, :16] = ids[:, 16:32]              # duplicate keys inside every batch
    kb = np.stack([np.frombuffer("".join("q%08d" % i for i in row).encode(), dtype=np.uint8) for row in ids])
    off = torch.from_numpy(np.arange(B + 1, dtype=np.int64) * 9).cuda()
    x = rng.standard_normal((n, B, D)).astype(np.float32)
    w = (rng.standard_normal((D, D)) / np.sqrt(D)).astype(np.float32)
    bias = rng.standard_normal(D).astype(np.float32)
    a = DispatchPipeline(max_size, B, D, D, model, w, bias, default_ttl=150.0)
    b = DispatchPipeline(max_size, B, D, D, model, w, bias, default_ttl=150.0)
    kd, xd = torch.from_numpy(kb).cuda(), torch.from_numpy(x).cuda()
    out = torch.empty((n, B, D), device="cuda")
    hit = torch.empty((n, B), dtype=torch.uint8, device="cuda")
    launches0, replays0 = a._lib.die_launch_count(), a._lib.die_replay_count()
    a.run_device_many(kd, off, xd, 5.0, 1.5, out, hit)
    torch.cuda.synchronize()
    # echo: three kernels per batch from a replayed graph; linear_tf32: one persistent kernel per group of 16 batches
    assert a._lib.die_launch_count() - launches0 >= (3 * n if model == "echo" else n // 16 + n % 16)
    assert a._lib.die_replay_count() - replays0 == n // 16
    for j in range(n):
        o, h = b.run_device(ops.KeyBatch(kd[j], off), xd[j], 5.0 + j * 1.5)
        assert torch.equal(h, hit[j]), j
        assert torch.equal(o, out[j]), j
    # the same queue from pinned host buffers: groups of 16 batches per copy and per replay
    c = DispatchPipeline(max_size, B, D, D, model, w, bias, default_ttl=150.0)
    hout = torch.empty((n, B, D)).pin_memory()
    hhit = torch.empty((n, B), dtype=torch.uint8).pin_memory()
    kbp = torch.from_numpy(np.concatenate([kb, np.zeros((n, 16), np.uint8)], axis=1)).pin_memory()
    replays0 = c._lib.die_replay_count()
    c.run_host_many(kbp, off.cpu().pin_memory(), torch.from_numpy(x).pin_memory(), 5.0, 1.5, hout, hhit)
    assert c._lib.die_replay_count() - replays0 == (n + 15) // 16 + 1  # the first 16 batches go as 4 + 12 (short ramp)
    assert torch.equal(hhit, hit.cpu()) and torch.equal(hout, out.cpu())
    sa, sb, sc = a.kv.stats(), b.kv.stats(), c.kv.stats()
    for k in ("hits", "misses", "evictions", "size"):
        assert sa[k] == sb[k] == sc[k], k
    assert sa["evictions"] > 0 and sa["hits"] > 0
    da, db = a.kv.dump(), b.kv.dump()
    for name in da:
        if name == "values":
            continue  # result-row handles come from a free list: which row a key gets is not part of the contract
        assert torch.equal(da[name], db[name]), name


def test_benchmarked_configuration_vs_oracle():
    """The configuration bench.py times (BASELINE.json configs[1]): kv_max_size 2**20, ids uniform in [0, 1e7),
    B = 4096, 128-dim payload, linear_tf32, queues submitted through die_pipeline_run_device_many with graph
    replays on -- 352 batches, i.e. past the point where the table is full and every batch evicts.  Hit flags per
    batch, counters and the final LRU order of all 2**20 entries are bit-exact against the oracle; outputs within
    2**-10 * sum|x w| + 1e-6.  The payload is a function of the request id, so a cached row and a recomputed row
    have the same expected value and the oracle only has to remember the id (not 2**20 result rows)."""
    from die.pipeline import DispatchPipeline
    B, D, kv_max, space, ttl, n_batches, chunk = 4096, 128, 1 << 20, 10 ** 7, 300.0, 352, 32
    rng = np.random.default_rng(2024)
    w = (rng.standard_normal((D, D)) / np.sqrt(D)).astype(np.float32)
    b = rng.standard_normal(D).astype(np.float32)
    base = rng.standard_normal((8192, D)).astype(np.float32)

    def payload_of(ids):
        return base[ids & 8191] * (1.0 + ((ids >> 13) % 1024).astype(np.float32)[:, None] / 1024.0)

    pipe = DispatchPipeline(kv_max, B, D, D, "linear_tf32", w, b, default_ttl=ttl, max_key_bytes=2 * B * 15)
    ref = OraclePipeline(1, kv_max, ttl, lambda rows: list(rows))   # oracle payload = the id; echo keeps it as the value
    off = torch.from_numpy(np.arange(B + 1, dtype=np.int64) * 15).cuda()
    out = torch.empty((chunk, B, D), dtype=torch.float32, device="cuda")
    hit = torch.empty((chunk, B), dtype=torch.uint8, device="cuda")
    now, dt = 1000.0, 0.01
    replays0 = pipe._lib.die_replay_count()
    for c0 in range(0, n_batches, chunk):
        ids = rng.integers(0, space, (chunk, B), dtype=np.int64)
        ids[:, :4] = ids[:, 4:8]                                     # duplicate keys inside every batch
        kb = np.frombuffer(b"".join(b"user-%010d" % i for i in ids.reshape(-1)), dtype=np.uint8).reshape(chunk, B * 15)
        x = payload_of(ids.reshape(-1)).reshape(chunk, B, D)
        pipe.run_device_many(torch.from_numpy(kb.copy()).cuda(), off, torch.from_numpy(x).cuda(), now, dt, out, hit)
        got_out, got_hit = out.cpu().numpy(), hit.cpu().numpy()
        for j in range(chunk):
            keys = ["user-%010d" % i for i in ids[j]]
            want = ref.process(keys, ids[j].tolist(), now + j * dt)
            assert got_hit[j].tolist() == want["hit"], c0 + j
            src = np.asarray(want["out"], dtype=np.int64)            # id whose payload produced each result
            xs = payload_of(src)
            err = np.abs(got_out[j].astype(np.float64) - models.linear_f64(xs, w, b))
            assert np.all(err <= 2.0 ** -10 * models.linear_abs_bound(xs, w) + 1e-6), c0 + j
        now += chunk * dt
    assert pipe._lib.die_replay_count() - replays0 == n_batches // 16  # the queues went through the captured replays
    st, rs = pipe.kv.stats(), ref.state()[0]
    assert (st["hits"], st["misses"], st["evictions"], st["size"]) == (rs["hits"], rs["misses"], rs["evictions"], rs["size"])
    assert rs["size"] == kv_max and rs["evictions"] > 20 * B and rs["hits"] > 0
    assert lru_digests(pipe) == [hashlib.md5(k.encode()).digest() for k in rs["lru"]]


@pytest.mark.parametrize("model", ["echo", "linear_tf32"])
def test_coordinator_on_device_path_vs_oracle(model):
    """die.Coordinator with caches larger than a batch: route (health gate + failover, router.py:140-221) -> bucket ->
    per-shard regions -> each shard's owner pass with its request count taken on the device, on its own stream ->
    gather.  Nothing of a batch comes back to the host.  Against OraclePipeline with the same unhealthy worker:
    shard ids, hit flags, per-shard counters and LRU order bit-exact; outputs exact (echo) / within the TF32 bound."""
    from die import Coordinator, ModelRegistry, Router
    n, B, D, max_size, ttl = 4, 512, 128, 1500, 9.0
    rng = np.random.default_rng(404)
    w = (rng.standard_normal((D, D)) / np.sqrt(D)).astype(np.float32)
    b = rng.standard_normal(D).astype(np.float32)
    reg = ModelRegistry()
    reg.register_model("m", "1", "/p", {}, {})
    router = Router(reg)
    for s in range(n):
        reg.add_shard("m", "1", s, f"w{s}")
        router.register_worker(f"w{s}", f"127.0.0.1:{9000 + s}")
        router.mark_worker_success(f"w{s}")
    for _ in range(3):
        router.mark_worker_failure("w2")   # its keys fail over to healthy[md5 % 3]

    if model == "echo":
        fn = lambda rows: [np.stack([r, np.zeros_like(r)]) for r in rows]  # noqa: E731
        co = Coordinator(router, "m", "1", max_size, B, D, model="echo", default_ttl=ttl)
    else:
        def fn(rows):
            if not rows:
                return []
            xs = np.stack(rows)
            y, bound = models.linear_f64(xs, w, b), models.linear_abs_bound(xs, w)
            return [np.stack([y[i], bound[i]]) for i in range(len(rows))]
        co = Coordinator(router, "m", "1", max_size, B, D, D, "linear_tf32", w, b, default_ttl=ttl)
    assert co.on_device
    ref = OraclePipeline(n, max_size, ttl, fn)
    for _ in range(3):
        ref.router.mark_worker_failure("w2")
    now = 10.0
    for step in range(24):
        now += float(rng.choice([0.5, 2.0, 4.0]))
        Bs = B if step % 3 else 300
        keys = ["user-%010d" % i for i in rng.integers(0, 9000, Bs)]
        x = rng.standard_normal((Bs, D)).astype(np.float32)
        out, hit, shard = co.process(keys, torch.from_numpy(x).cuda(), now)
        want = ref.process(keys, [x[i] for i in range(Bs)], now)
        assert shard.tolist() == want["shard"] and 2 not in want["shard"], step
        assert hit.tolist() == want["hit"], step
        wv = np.stack(want["out"])
        got = out.cpu().numpy()
        if model == "echo":
            assert np.array_equal(got, wv[:, 0].astype(np.float32)), step
        else:
            assert np.all(np.abs(got.astype(np.float64) - wv[:, 0]) <= 2.0 ** -10 * wv[:, 1] + 1e-6), step
    for s in range(n):
        st, ws = co.pipes[s].kv.stats(), ref.state()[s]
        assert (st["hits"], st["misses"], st["evictions"], st["size"]) == (ws["hits"], ws["misses"], ws["evictions"], ws["size"])
        assert lru_digests(co.pipes[s]) == [hashlib.md5(k.encode()).digest() for k in ws["lru"]]
    assert sum(s_["evictions"] for s_ in ref.state()) > 0 and sum(s_["hits"] for s_ in ref.state()) > 0
