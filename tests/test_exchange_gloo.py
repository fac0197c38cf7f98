"""world_size-2 test of the sharded dispatch bookkeeping (die/exchange.py) on CPU with the gloo backend.

The GPU stages are replaced by CPU stand-ins built from the oracle (this is the test's own scaffolding: the
product has no CPU path), so what is exercised is exactly the N>1 logic: split sizes, the dispatch and
return all-to-alls, owners applying sources in rank order, and the scatter back to request order.
Both ranks compare their results with oracle.pipeline.OraclePipeline.process_exchange."""
import hashlib
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class CpuStages:
    """Oracle-backed stand-ins with the interface of die.exchange.GpuStages."""

    def __init__(self, world, max_size, ttl, key_of_digest):
        from oracle import ref_port as rp
        self.world, self.key_of = world, key_of_digest
        self.now = 0.0
        self.cache = rp.PortKVCache(max_size, ttl, clock=lambda: self.now)
        self.dim_out = None

    def route(self, keys):
        from oracle import ref_port as rp
        shard = torch.tensor([rp.shard_index(k, self.world) for k in keys], dtype=torch.int32)
        raw = b"".join(hashlib.md5(k.encode()).digest() for k in keys)
        return shard, torch.from_numpy(np.frombuffer(raw, dtype=np.int32).reshape(-1, 4).copy())

    def bucket(self, shard):
        s = shard.to(torch.int64).clone()
        s[s < 0] = self.world
        perm = torch.argsort(s, stable=True).to(torch.int32)
        counts = torch.bincount(s, minlength=self.world + 1)
        offsets = torch.zeros(self.world + 2, dtype=torch.int64)
        offsets[1:] = torch.cumsum(counts, 0)
        pos = torch.empty_like(perm)
        pos[perm.long()] = torch.arange(perm.numel(), dtype=torch.int32)
        return offsets, perm, pos

    def gather_rows(self, rows, perm):
        return rows[perm.long()].contiguous()

    def scatter_rows(self, rows, perm, out):
        out[perm.long()] = rows
        return out

    def empty(self, shape, dtype):
        return torch.empty(shape, dtype=dtype)

    def owner(self, digests, payload, now):
        self.now = now
        n = digests.shape[0]
        out = torch.empty_like(payload)
        hit = torch.zeros(n, dtype=torch.uint8)
        raw = digests.numpy().tobytes()
        keys = [self.key_of[raw[16 * i:16 * i + 16]] for i in range(n)]
        miss = []
        for i, k in enumerate(keys):
            v = self.cache.get(k)
            if v is None:
                miss.append(i)
            else:
                hit[i] = 1
                out[i] = v
        for i in miss:              # echo model
            out[i] = payload[i]
        for i in miss:
            self.cache.set(keys[i], out[i].clone())
        return out, hit


def _worker(rank, world, port, ret):
    sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-inference-engine_b200")]
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from die.exchange import ShardedDispatcher
        from oracle.pipeline import OraclePipeline
        B, D, steps, max_size, ttl = 48, 4, 12, 20, 6.0
        rng = np.random.default_rng(123)  # same stream on both ranks: each builds everybody's inputs
        all_keys = ["user-%010d" % i for i in range(90)]
        key_of = {hashlib.md5(k.encode()).digest(): k for k in all_keys}
        disp = ShardedDispatcher(world, rank, stages=CpuStages(world, max_size, ttl, key_of))
        ref = OraclePipeline(world, max_size, ttl, model=lambda rows: list(rows))
        now = 10.0
        for _ in range(steps):
            now += float(rng.choice([0.5, 2.0, 5.0]))
            per_rank = []
            for _r in range(world):
                keys = [all_keys[i] for i in rng.integers(0, len(all_keys), B)]
                pay = rng.integers(-50, 50, (B, D)).astype(np.float32)
                per_rank.append((keys, pay))
            keys, pay = per_rank[rank]
            out, hit, shard = disp.step(keys, torch.from_numpy(pay), now)
            want = ref.process_exchange([(k, [row for row in p]) for k, p in per_rank], now)[rank]
            assert shard.tolist() == want["shard"]
            assert hit.tolist() == want["hit"]
            assert np.array_equal(out.numpy(), np.stack(want["out"]))
        mine = ref.state()[rank]
        st = disp.stages.cache.stats()
        assert (st["hits"], st["misses"], st["evictions"], st["size"]) == (mine["hits"], mine["misses"], mine["evictions"], mine["size"])
        assert disp.stages.cache.lru_order() == mine["lru"]
        ret[rank] = "ok"
    finally:
        dist.destroy_process_group()


def test_two_rank_exchange_matches_oracle_over_gloo():
    world = 2
    port = 29500 + (os.getpid() % 2000)
    with mp.Manager() as m:
        ret = m.dict()
        mp.spawn(_worker, args=(world, port, ret), nprocs=world, join=True)
        assert dict(ret) == {0: "ok", 1: "ok"}
