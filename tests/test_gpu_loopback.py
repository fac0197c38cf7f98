"""GPU parity of the sharded (N > 1) step on a ONE-GPU box: all ranks of the peer exchange run in this process on
one device (die.exchange.LoopbackPeerGroup), through the same die_xchg_step / die_xchg_step_many entry points
and the same csrc/exchange.cu kernels that run one rank per GPU over NVLink.  Compared with
oracle.pipeline.OraclePipeline.process_exchange: owner = md5 mod N (model_registry.py:159-161), owners apply
the sources in rank order, request order inside a source (kvstore.py:104-164 semantics per owner).  Shard ids,
hit flags, counters and every owner's LRU order are bit-exact; echo outputs bit-exact; TF32 outputs within
2**-10 * sum|x w| + 1e-6 (oracle/models.py)."""
import hashlib
import os

import numpy as np
import pytest
import torch

from oracle import models
from oracle import ref_port as rp
from oracle.pipeline import OraclePipeline

pytestmark = pytest.mark.gpu


def lru_digests(pipe):
    d = pipe.kv.dump()
    return [bytes(x) for x in d["digests"].numpy().view(np.uint8).reshape(-1, 16)]


def make_model(model, D, rng):
    w = (rng.standard_normal((D, D)) / np.sqrt(D)).astype(np.float32)
    b = rng.standard_normal(D).astype(np.float32)
    if model == "echo":
        return None, None, (lambda rows: [np.stack([r, np.zeros_like(r)]) for r in rows])
    if model == "mlp_bf16":
        H = 2 * D
        w1 = models.to_bf16((rng.standard_normal((H, D)) / np.sqrt(D)).astype(np.float32))
        w2 = models.to_bf16((rng.standard_normal((D, H)) / np.sqrt(H)).astype(np.float32))
        b1, b2 = rng.standard_normal(H).astype(np.float32), rng.standard_normal(D).astype(np.float32)

        def mlp(rows):  # value = (result, per-row tolerance of oracle/models.py: 2**-7 |y| + 1e-2 max|y|)
            if not rows:
                return []
            y = models.mlp_bf16(models.to_bf16(np.stack(rows)), w1, b1, w2, b2)
            tol = 2.0 ** -7 * np.abs(y) + 1e-2 * np.abs(y).max(axis=1, keepdims=True)
            return [np.stack([y[i], tol[i]]) for i in range(len(rows))]
        return (w1, w2), (b1, b2), mlp

    def fn(rows):  # value = (result, error scale sum|x w|): cached rows keep the bound of the payload they came from
        if not rows:
            return []
        xs = np.stack(rows)
        y, bound = models.linear_f64(xs, w, b), models.linear_abs_bound(xs, w)
        return [np.stack([y[i], bound[i]]) for i in range(len(rows))]
    return w, b, fn


def check_outputs(model, out, want_rows, where):
    want = np.stack(want_rows)
    got = out.cpu().numpy()
    if model == "echo":
        assert np.array_equal(got, want[:, 0].astype(np.float32)), where
    elif model == "mlp_bf16":
        assert np.all(np.abs(got.astype(np.float64) - want[:, 0]) <= want[:, 1]), where
    else:
        rel = 2.0 ** -10 if model == "linear_tf32" else 1e-5
        assert np.all(np.abs(got.astype(np.float64) - want[:, 0]) <= rel * want[:, 1] + 1e-6), where


def check_owner_state(group, ref):
    for r in range(group.world):
        st, want = group.pipes[r].kv.stats(), ref.state()[r]
        assert (st["hits"], st["misses"], st["evictions"], st["size"]) == \
               (want["hits"], want["misses"], want["evictions"], want["size"]), r
        assert lru_digests(group.pipes[r]) == [hashlib.md5(k.encode()).digest() for k in want["lru"]], r


@pytest.mark.parametrize("world,model", [(2, "echo"), (2, "linear_tf32"), (4, "linear"), (2, "mlp_bf16")])
def test_loopback_sharded_steps_match_oracle(world, model):
    """Stream-launched steps (die_xchg_step), small owner tables: evictions, TTL expiry, duplicate keys inside
    and across the ranks' batches."""
    from die import ops
    from die.exchange import LoopbackPeerGroup
    B, D, max_size, ttl, n_ids, steps = 256, (256 if model == "mlp_bf16" else 128), 700, 6.0, 1250 * world, 14
    rng = np.random.default_rng(100 + world)
    w, b, fn = make_model(model, D, rng)
    group = LoopbackPeerGroup(world, max_size, B, D, w, b, ttl, model=model)
    ref = OraclePipeline(world, max_size, ttl, fn)
    now = 3.0
    for step in range(steps):
        now += float(rng.choice([0.5, 2.0, 4.0]))
        per_rank = []
        for _r in range(world):
            ids = rng.integers(0, n_ids, B)
            ids[:8] = ids[8:16]  # duplicates inside a batch
            per_rank.append((["user-%010d" % i for i in ids], rng.standard_normal((B, D)).astype(np.float32)))
        res = group.step([ops.KeyBatch.from_strings(k) for k, _ in per_rank],
                         [torch.from_numpy(p).cuda() for _, p in per_rank], now)
        want = ref.process_exchange([(k, [r for r in p]) for k, p in per_rank], now)
        for r in range(world):
            out, hit, shard = res[r]
            assert shard.tolist() == want[r]["shard"], (step, r)
            assert hit.tolist() == want[r]["hit"], (step, r)
            check_outputs(model, out, want[r]["out"], (step, r))
        group.maintain()  # housekeeping between steps: inside a step it would wait for peers not yet enqueued
    check_owner_state(group, ref)
    assert sum(s["evictions"] for s in ref.state()) > 0 and sum(s["hits"] for s in ref.state()) > 0


@pytest.mark.parametrize("owner_group", ["0", "1"])
def test_loopback_replayed_queue_matches_oracle(monkeypatch, owner_group):
    """die_xchg_step_many over many calls, until the owners' tables evict.  owner_group=1 (the default): the
    software-pipelined replays -- per graph launch the ingest of group g, ONE persistent owner launch over group g-1
    and the way back of group g-2, on three thirds of the buffer sets, barrier epochs from the device descriptor.
    owner_group=0: whole steps captured into one graph per group, one owner launch per step."""
    from die import ops
    from die.exchange import LoopbackPeerGroup
    monkeypatch.setenv("DIE_GRAPH_STEPS", "4")
    monkeypatch.setenv("DIE_XCHG_OWNER_GROUP", owner_group)
    world, B, D, max_size, ttl, n_ids, K, groups = 2, 128, 128, 8192, 40.0, 60000, 4, 40
    rng = np.random.default_rng(77)
    model = "linear_tf32"
    w, b, fn = make_model(model, D, rng)
    group = LoopbackPeerGroup(world, max_size, B, D, w, b, ttl, model=model)
    ref = OraclePipeline(world, max_size, ttl, fn)
    now, dt = 5.0, 0.25
    replays0 = group.lib.die_replay_count()
    for g in range(groups):
        batches = [[(["user-%010d" % i for i in rng.integers(0, n_ids, B)],
                     rng.standard_normal((B, D)).astype(np.float32)) for _r in range(world)] for _ in range(K)]
        kbs = [[ops.KeyBatch.from_strings(bt[r][0]) for bt in batches] for r in range(world)]
        res = group.step_many([torch.stack([k.key_bytes for k in kbs[r]]) for r in range(world)],
                              [torch.stack([k.key_off for k in kbs[r]]) for r in range(world)],
                              [torch.from_numpy(np.stack([bt[r][1] for bt in batches])).cuda() for r in range(world)],
                              now, dt)
        for j, bt in enumerate(batches):
            want = ref.process_exchange([(k, [r for r in p]) for k, p in bt], now + j * dt)
            for r in range(world):
                assert res[r][1][j].tolist() == want[r]["hit"], (g, j, r)
                check_outputs(model, res[r][0][j], want[r]["out"], (g, j, r))
        now += K * dt
        group.maintain()
    assert group.lib.die_replay_count() - replays0 == world * groups
    check_owner_state(group, ref)
    assert all(s["evictions"] > 0 for s in ref.state())


@pytest.mark.parametrize("failover", [True, False])
def test_loopback_sharded_steps_follow_the_router_health_gate(failover):
    """Router.route_request's health gate and _find_alternative_shard (router.py:165-221) on the fused sharded step
    (die_xchg_set_health): one of four owners is marked unhealthy after a few steps and recovers later.  With failover
    its keys go to healthy_idx[md5 mod n_healthy]; without, those requests have no owner (shard -1, zero row, no hit).
    Single steps and a replayed queue; shard ids, hit flags, results, owner tables against the oracle's PortRouter."""
    from die import ops
    from die.exchange import LoopbackPeerGroup
    world, B, D, max_size, ttl, n_ids = 4, 256, 128, 900, 50.0, 3000
    rng = np.random.default_rng(5 if failover else 6)
    model = "linear"
    w, b, fn = make_model(model, D, rng)
    group = LoopbackPeerGroup(world, max_size, B, D, w, b, ttl, model=model)
    ref = OraclePipeline(world, max_size, ttl, fn, failover_enabled=failover)

    def healthy_flags():
        return [ref.router.workers["w%d" % s][0] == rp.HEALTHY for s in range(world)]

    def step(now):
        per_rank = [(["user-%010d" % i for i in rng.integers(0, n_ids, B)],
                     rng.standard_normal((B, D)).astype(np.float32)) for _ in range(world)]
        res = group.step([ops.KeyBatch.from_strings(k) for k, _ in per_rank],
                         [torch.from_numpy(p).cuda() for _, p in per_rank], now)
        want = ref.process_exchange([(k, [r for r in p]) for k, p in per_rank], now)
        for r in range(world):
            out, hit, shard = res[r]
            assert shard.tolist() == want[r]["shard"], r
            assert hit.tolist() == want[r]["hit"], r
            rows = [np.stack([np.zeros(D), np.zeros(D)]) if o is None else o for o in want[r]["out"]]
            check_outputs(model, out, rows, r)
        group.maintain()   # table housekeeping between steps: inside one it would synchronise the device all ranks share
        return sum(1 for r in range(world) for s in want[r]["shard"] if s < 0)

    now = 1.0
    for _ in range(3):
        now += 1.0
        assert step(now) == 0
    for _ in range(3):
        ref.router.mark_worker_failure("w2")
    flags = healthy_flags()
    assert flags == [True, True, False, True]
    group.set_health([1 if f else 0 for f in flags], [i for i, f in enumerate(flags) if f], failover)
    dropped = 0
    for _ in range(4):
        now += 1.0
        dropped += step(now)
    assert (dropped == 0) == failover
    ref.router.mark_worker_success("w2")
    group.set_health(None)
    for _ in range(3):
        now += 1.0
        assert step(now) == 0
    check_owner_state(group, ref)
