"""GPU: stable bucket-by-shard kernel, and (when the box has >= 2 GPUs) the NCCL sharded dispatch vs the oracle."""
import os
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("B,n", [(4096, 8), (1, 1), (10000, 3), (257, 64), (5000, 2)])
def test_bucket_by_shard_is_a_stable_partition(B, n):
    from die import ops
    rng = np.random.default_rng(B + n)
    shard = rng.integers(-1, n, B).astype(np.int32)
    offsets, perm, pos = ops.bucket_by_shard(torch.from_numpy(shard).cuda(), n)
    key = np.where(shard < 0, n, shard)
    want_perm = np.argsort(key, kind="stable")
    assert np.array_equal(perm.cpu().numpy(), want_perm)
    assert np.array_equal(pos.cpu().numpy()[want_perm], np.arange(B))
    counts = np.bincount(key, minlength=n + 1)
    assert offsets.cpu().tolist() == [0] + np.cumsum(counts).tolist()


def _worker(rank, world, port, ret, mode="nccl"):
    sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-inference-engine_b200")]
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from die import ops
        from die.exchange import PeerDispatcher, ShardedDispatcher
        from oracle import models
        from oracle.pipeline import OraclePipeline
        B, D, steps, max_size, ttl, n_ids = 1024, 128, 8, 3000, 7.0, 5000
        model = "linear_tf32" if mode == "peer_queue_tf32" else "linear"
        if mode.startswith("peer_queue"):  # a table large enough for captured replays of 8 steps, filled to eviction
            steps, max_size, ttl, n_ids = 93, 70000, 60.0, 400000
            os.environ["DIE_GRAPH_STEPS"] = "8"
        rng = np.random.default_rng(5)
        w = (rng.standard_normal((D, D)) / np.sqrt(D)).astype(np.float32)
        b = rng.standard_normal(D).astype(np.float32)
        if mode == "nccl":
            disp = ShardedDispatcher(world, rank, max_size, B, D, w, b, ttl, model="linear")
        else:
            disp = PeerDispatcher(world, rank, max_size, B, D, w, b, ttl, model=model, single_call=mode != "peer_phases")

        def fn(rows):  # value = (result, error scale sum|x w|): a cached row keeps the bound of its payload
            if not rows:
                return []
            xs = np.stack(rows)
            y, bound = models.linear_f64(xs, w, b), models.linear_abs_bound(xs, w)
            return [np.stack([y[i], bound[i]]) for i in range(len(rows))]

        def close(got, want_rows):
            want = np.stack(want_rows)
            rel = 2.0 ** -10 if model == "linear_tf32" else 1e-5
            return np.all(np.abs(got.cpu().numpy().astype(np.float64) - want[:, 0]) <= rel * want[:, 1] + 1e-6)

        ref = OraclePipeline(world, max_size, ttl, fn)
        now = 3.0
        if mode.startswith("peer_queue"):
            # peer_queue: one graph per group of whole steps (the fp32 owner pipeline has no persistent form);
            # peer_queue_tf32: the software-pipelined replays with one persistent owner launch per group.
            # a queue of batches through the two-lane step: neighbouring steps overlap on the GPU, the KV
            # table must still see them in step order (evictions and TTL expiry depend on it)
            dt = 0.9
            batches = []
            for _ in range(steps):
                batches.append([(["user-%010d" % i for i in rng.integers(0, n_ids, B)],
                                 rng.standard_normal((B, D)).astype(np.float32)) for _r in range(world)])
            kb = [ops.KeyBatch.from_strings(bt[rank][0]) for bt in batches]
            key_bytes = torch.stack([k.key_bytes for k in kb])
            key_off = torch.stack([k.key_off for k in kb])
            pay = torch.from_numpy(np.stack([bt[rank][1] for bt in batches])).cuda()
            out = torch.empty((steps, B, D), dtype=torch.float32, device="cuda")
            hit = torch.empty((steps, B), dtype=torch.uint8, device="cuda")
            replays0 = disp.lib.die_replay_count()
            disp.step_many(key_bytes, key_off, pay, now, dt, out, hit)
            torch.cuda.synchronize()
            assert disp.lib.die_replay_count() - replays0 == steps // 8
            for j, bt in enumerate(batches):
                want = ref.process_exchange([(k, [r for r in p]) for k, p in bt], now + j * dt)[rank]
                assert hit[j].tolist() == want["hit"], j
                assert close(out[j], want["out"]), j
            assert ref.state()[rank]["evictions"] > 0
            steps = 0
        for _ in range(steps):
            now += float(rng.choice([0.5, 2.0, 4.0]))
            per_rank = []
            for _r in range(world):
                keys = ["user-%010d" % i for i in rng.integers(0, 5000, B)]
                per_rank.append((keys, rng.standard_normal((B, D)).astype(np.float32)))
            keys, pay = per_rank[rank]
            out, hit, shard = disp.step(ops.KeyBatch.from_strings(keys), torch.from_numpy(pay).cuda(), now)
            want = ref.process_exchange([(k, [r for r in p]) for k, p in per_rank], now)[rank]
            assert shard.tolist() == want["shard"] and hit.tolist() == want["hit"]
            assert close(out, want["out"])
        st, mine = disp.pipe.kv.stats(), ref.state()[rank]
        assert (st["hits"], st["misses"], st["evictions"], st["size"]) == (mine["hits"], mine["misses"], mine["evictions"], mine["size"])
        if mode != "nccl":
            disp.check()
        ret[rank] = "ok"
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (run with gpurun --gpus 2)")
@pytest.mark.parametrize("mode", ["nccl", "peer_phases", "peer_one_call", "peer_queue", "peer_queue_tf32"])
def test_two_gpu_sharded_dispatch_matches_oracle(mode):
    import torch.multiprocessing as mp
    world = 2
    port = 29600 + (os.getpid() % 2000) + len(mode)
    with mp.Manager() as m:
        ret = m.dict()
        mp.spawn(_worker, args=(world, port, ret, mode), nprocs=world, join=True)
        assert dict(ret) == {0: "ok", 1: "ok"}


def _ragged_worker(rank, world, port, ret, peer=False):
    """BASELINE.json configs[2]: ragged 16..2048-element payloads, max_batch 256, hash-sharded with the all-to-all
    dispatch; echo model (the reference's).  Batcher pack -> sharded dispatch -> unpack, vs the oracle."""
    sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-inference-engine_b200")]
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        from die import ops
        from die.batcher import PackedBatch, pack_requests, unpack_results
        from die.exchange import PeerDispatcher, ShardedDispatcher
        from oracle.pipeline import OraclePipeline
        B, Lpad, steps, max_size = 256, 2048, 6, 400
        if peer:
            max_size = 600  # the fused owner pipeline needs max_size >= its batch capacity (2 x 256)
        rng = np.random.default_rng(17)
        if peer:   # 8 KB rows through the NVLink peer-store step
            disp = PeerDispatcher(world, rank, max_size, B, Lpad, None, None, 50.0, model="echo")
        else:
            disp = ShardedDispatcher(world, rank, max_size, B, Lpad, None, None, 50.0, model="echo")
        ref = OraclePipeline(world, max_size, 50.0, lambda rows: list(rows))
        now = 1.0
        for _ in range(steps):
            now += 1.0
            per_rank = []
            for _r in range(world):
                keys = ["sess-%06d" % i for i in rng.integers(0, 700, B)]
                pays = [rng.standard_normal(int(n)).astype(np.float32) for n in rng.integers(16, 2049, B)]
                per_rank.append((keys, pays))
            keys, pays = per_rank[rank]
            packed = pack_requests(pays, Lpad=Lpad)                     # K4: ragged -> [256, 2048] zero padded
            out, hit, shard = disp.step(ops.KeyBatch.from_strings(keys), packed.data, now)
            # a hit returns the cached result of an EARLIER request with the same key: its length is the cached row's
            want = ref.process_exchange(per_rank, now)[rank]
            lens = [len(w) for w in want["out"]]
            got = unpack_results(PackedBatch(out, None, lens))          # K5: padded rows -> ragged results
            assert shard.tolist() == want["shard"] and hit.tolist() == want["hit"]
            for g, w_ in zip(got, want["out"]):
                assert np.array_equal(g.cpu().numpy(), w_)
        if peer:
            disp.check()
        ret[rank] = "ok"
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs 2 GPUs (run with gpurun --gpus 2)")
@pytest.mark.parametrize("peer", [False, True], ids=["nccl", "peer"])
def test_two_gpu_ragged_c3_dispatch_matches_oracle(peer):
    import torch.multiprocessing as mp
    port = 29700 + (os.getpid() % 2000) + (7 if peer else 0)
    with mp.Manager() as m:
        ret = m.dict()
        mp.spawn(_ragged_worker, args=(2, port, ret, peer), nprocs=2, join=True)
        assert dict(ret) == {0: "ok", 1: "ok"}
