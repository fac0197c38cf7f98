"""GPU parity: K2/K3 KV store through the C-ABI.  Every check is bit-exact: results per op, hit/miss,
LRU order, counters, entry metadata -- against traces recorded from the reference's KVCache and
against the CPU oracle on seeded random batches."""
import hashlib
import math

import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import ref_port as rp

pytestmark = pytest.mark.gpu


def md5_words(keys):
    raw = b"".join(hashlib.md5(k.encode()).digest() for k in keys)
    return torch.from_numpy(np.frombuffer(raw, dtype=np.int32).reshape(-1, 4).copy()).cuda()


# ------------------------------------------------------------------ reference traces on the drop-in class
def replay_on_drop_in(t):
    from die.kvstore import KVCache
    clock = [1000.0]
    kv = KVCache(t["max_size"], "lru", t["default_ttl"], clock=lambda: clock[0])
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
                r = op[1] in kv
            elif name == "len":
                r = len(kv)
            elif name == "clear":
                r = kv.clear()
            elif name == "stats":
                r = kv.get_stats()
        except StopIteration:
            r = "!StopIteration"
        res.append(r)
    return kv, res


@pytest.mark.parametrize("name", sorted(load_golden("kv_traces.json").keys()))
def test_reference_trace_on_drop_in_kvcache(name):
    t = load_golden("kv_traces.json")[name]
    kv, res = replay_on_drop_in(t)
    assert res == t["results"]
    assert kv.get_stats() == t["final"]["stats"]
    assert list(kv._access_order) == t["final"]["lru"]
    got = {k: [e.value, e.created_at, e.last_accessed, e.access_count, e.ttl] for k, e in kv.cache.items()}
    assert got == t["final"]["entries"]
    assert kv.hits == t["final"]["stats"]["hits"] and kv.evictions == t["final"]["stats"]["evictions"]
    assert kv.max_size == t["max_size"] and kv.eviction_policy == "lru"


def replay_policy_trace_on_drop_in(t):
    from die.kvstore import KVCache
    clock = [1000.0]
    kv = KVCache(t["max_size"], t["policy"], t["default_ttl"], clock=lambda: clock[0])
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
                r = op[1] in kv
            elif name == "len":
                r = len(kv)
            elif name == "clear":
                r = kv.clear()
            elif name == "stats":
                r = kv.get_stats()
        except StopIteration:
            r = "!StopIteration"
        except ValueError:
            r = "!ValueError"
        res.append(r)
    return kv, res


@pytest.mark.parametrize("name", sorted(load_golden("kv_policy_traces.json").keys()))
def test_reference_lfu_fifo_trace_on_drop_in_kvcache(name):
    """LFU and FIFO (kvstore.py:67-69,79-80,94-102): traces recorded from the reference's KVCache, replayed on the
    device table -- every result, the counters, every entry's metadata, the LFU use counts in the order the keys
    were first counted, the FIFO insertion order.  Bit-exact, including the ValueError / StopIteration cases."""
    t = load_golden("kv_policy_traces.json")[name]
    kv, res = replay_policy_trace_on_drop_in(t)
    assert res == t["results"]
    assert kv.get_stats() == t["final"]["stats"]
    got = {k: [e.value, e.created_at, e.last_accessed, e.access_count, e.ttl] for k, e in kv.cache.items()}
    assert got == t["final"]["entries"]
    if t["policy"] == "lfu":
        assert [[k, c] for k, c in kv._access_count.items()] == t["final"]["lfu"]
    elif not any(r in ("!StopIteration",) for r in t["results"] if isinstance(r, str)):
        assert list(kv.cache) == t["final"]["cache_order"]  # FIFO: the dict's insertion order is the eviction order


def test_unknown_policy_is_refused():
    from die.kvstore import KVCache
    with pytest.raises(ValueError):
        KVCache(4, "mru")


@pytest.mark.parametrize("policy", ["lru", "lfu", "fifo"])
def test_snapshot_restore_round_trip(policy, tmp_path):
    """save -> fresh cache -> load: the restored table answers, counts and evicts exactly like the original (the
    twin keeps running beside it as the reference), whatever the policy."""
    from die.kvstore import KVCache
    rng = np.random.default_rng(17)
    clock = [50.0]
    a = KVCache(40, policy, 30.0, clock=lambda: clock[0])

    def drive(kv, seed, n):
        r = np.random.default_rng(seed)
        out = []
        for _ in range(n):
            clock_step = float(r.choice([0.0, 0.5, 2.0]))
            clock[0] += clock_step
            k = "k%d" % int(r.integers(0, 90))
            if r.random() < 0.5:
                out.append(kv.get(k, "DFLT"))
            else:
                kv.set(k, int(r.integers(0, 10 ** 6)), ttl=None if r.random() < 0.7 else 5.0)
        return out

    drive(a, 1, 300)
    t_saved = clock[0]
    path = str(tmp_path / "cache.img")
    a.save(path)
    b = KVCache(40, policy, 30.0, clock=lambda: clock[0])
    b.load(path)
    assert b.get_stats() == a.get_stats() and list(b.cache) == list(a.cache)
    res_a = drive(a, 2, 300)
    state_a = (a.get_stats(), {k: (e.value, e.access_count) for k, e in a.cache.items()})
    clock[0] = t_saved
    res_b = drive(b, 2, 300)
    assert res_a == res_b
    assert state_a == (b.get_stats(), {k: (e.value, e.access_count) for k, e in b.cache.items()})
    assert a.get_stats()["evictions"] > 0
    with pytest.raises(Exception):
        KVCache(41, policy).load(path)  # another shape: refused
    del rng


# ------------------------------------------------------------------ batched device API vs the oracle
class PortMirror:
    """Applies the same batches to the CPU oracle, sequentially in request order."""

    def __init__(self, max_size, default_ttl):
        self.now = 0.0
        self.kv = rp.PortKVCache(max_size, default_ttl, clock=lambda: self.now)

    def get_batch(self, keys, now):
        self.now = now
        out = []
        for k in keys:
            v = self.kv.get(k, None)
            out.append(v)
        return out

    def set_batch(self, keys, values, ttls, now):
        self.now = now
        for k, v, t in zip(keys, values, ttls):
            self.kv.set(k, v, ttl=t)


def compare_state(dkv, port, digest_of):
    st = dkv.stats()
    ps = port.kv.stats()
    assert (st["hits"], st["misses"], st["evictions"], st["size"]) == (ps["hits"], ps["misses"], ps["evictions"], ps["size"])
    d = dkv.dump()
    order = [bytes(x) for x in d["digests"].numpy().view(np.uint8).reshape(-1, 16)]
    stamped = d["stamped"].numpy().astype(bool)
    want = [digest_of(k) for k in port.kv.lru_order()]
    assert [o for o, s in zip(order, stamped) if s] == want
    vals = dict(zip(order, d["values"].tolist()))
    created = dict(zip(order, d["created_at"].tolist()))
    cnt = dict(zip(order, d["access_count"].tolist()))
    ttl = dict(zip(order, d["ttl"].tolist()))
    for k, e in port.kv.cache.items():
        dg = digest_of(k)
        assert vals[dg] == e.value and created[dg] == e.created_at and cnt[dg] == e.access_count
        assert (e.ttl is None and math.isnan(ttl[dg])) or ttl[dg] == e.ttl


@pytest.mark.parametrize("cfg", [
    # max_size, default_ttl, key space, batches, max batch, ttl choices
    (64, None, 200, 60, 48, [None]),
    (500, 20.0, 900, 40, 700, [None, None, 5.0, 0.0]),
    (4000, None, 6000, 25, 3000, [None, 30.0]),
    (37, 3.0, 60, 80, 90, [None, 1.0, 8.0]),       # batches larger than the cache: sequential path
    (3, None, 10, 60, 7, [None]),
    (1, None, 4, 40, 5, [None, 1.0]),
])
def test_random_batches_match_oracle(cfg):
    from die.kvstore import DeviceKV
    max_size, dttl, space, n_batches, max_b, ttl_choices = cfg
    rng = np.random.default_rng(max_size * 7 + space)
    dkv = DeviceKV(max_size, dttl)
    port = PortMirror(max_size, dttl)
    cache_digest = {}

    def digest_of(k):
        if k not in cache_digest:
            cache_digest[k] = hashlib.md5(k.encode()).digest()
        return cache_digest[k]

    now, next_val = 100.0, 1
    for b in range(n_batches):
        now += float(rng.choice([0.0, 0.5, 2.0, 6.0]))
        B = int(rng.integers(1, max_b + 1))
        keys = ["k%d" % i for i in rng.integers(0, space, B)]
        dig = md5_words(keys)
        if rng.random() < 0.5:
            hit, value = dkv.get_batch(dig, now)
            want = port.get_batch(keys, now)
            assert [int(h == 1) for h in hit.tolist()] == [int(v is not None) for v in want], b
            got_vals = [v if h == 1 else None for h, v in zip(hit.tolist(), value.tolist())]
            assert got_vals == want, b
        else:
            values = list(range(next_val, next_val + B))
            next_val += B
            ttls = [ttl_choices[i] for i in rng.integers(0, len(ttl_choices), B)]
            tt = torch.tensor([math.nan if t is None else t for t in ttls], dtype=torch.float64, device="cuda")
            dkv.set_batch(dig, torch.tensor(values, dtype=torch.int64, device="cuda"), now, tt)
            port.set_batch(keys, values, ttls, now)
            status, _, _ = dkv.sync_status()
            assert status == 0
        if b % 7 == 3:
            dkv.maintain(force=True)  # LRU-log renumbering + table rebuild must not change anything
        compare_state(dkv, port, digest_of)


def test_fast_path_is_taken_when_cache_is_large_and_slow_path_when_forced():
    from die.kvstore import DeviceKV
    dkv = DeviceKV(1000, None)
    keys = ["a%d" % i for i in range(3000)]
    for i in range(0, 3000, 500):
        dkv.set_batch(md5_words(keys[i:i + 500]), torch.arange(i, i + 500, dtype=torch.int64, device="cuda"), 1.0)
    status, _, slow = dkv.sync_status()
    assert status == 0 and slow == 0          # evictions served by the parallel path
    st = dkv.stats()
    assert st["size"] == 1000 and st["evictions"] == 2000
    # re-set the LRU head inside an evicting batch: Appendix A.2 #1 order dependence -> exact replay
    dkv.set_batch(md5_words(["new-key", keys[2000]]), torch.tensor([7, 8], dtype=torch.int64, device="cuda"), 2.0)
    status, _, slow = dkv.sync_status()
    assert status == 0 and slow == 1
    assert dkv.stats()["evictions"] == 2002   # head evicted, then re-inserted as new -> a second eviction


def test_pipeline_style_fill_at_scale_matches_oracle():
    """get -> set(misses) rounds on a 200k-entry table, 32768 ops per batch."""
    from die.kvstore import DeviceKV
    max_size, B, rounds, space = 200_000, 32768, 14, 400_000
    rng = np.random.default_rng(11)
    dkv = DeviceKV(max_size, 50.0)
    port = PortMirror(max_size, 50.0)
    now = 10.0
    for r in range(rounds):
        now += 4.0
        ids = rng.integers(0, space, B)
        keys = ["user-%010d" % i for i in ids]
        dig = md5_words(keys)
        hit, value = dkv.get_batch(dig, now)
        want = port.get_batch(keys, now)
        hl = hit.cpu().numpy()
        assert np.array_equal(hl == 1, np.array([v is not None for v in want])), r
        miss = np.nonzero(hl != 1)[0]
        mkeys = [keys[i] for i in miss]
        vals = [int(ids[i]) + 1 for i in miss]
        dkv.set_batch(dig[torch.from_numpy(miss).cuda()], torch.tensor(vals, dtype=torch.int64, device="cuda"), now)
        port.set_batch(mkeys, vals, [None] * len(vals), now)
    status, _, slow = dkv.sync_status()
    st, ps = dkv.stats(), port.kv.stats()
    assert status == 0 and slow == 0
    assert (st["hits"], st["misses"], st["evictions"], st["size"]) == (ps["hits"], ps["misses"], ps["evictions"], ps["size"])
    d = dkv.dump()
    order = [bytes(x) for x in d["digests"].numpy().view(np.uint8).reshape(-1, 16)]
    assert order == [hashlib.md5(k.encode()).digest() for k in port.kv.lru_order()]


def test_contains_delete_len_clear_batched():
    from die.kvstore import DeviceKV
    dkv = DeviceKV(100, None)
    keys = ["c%d" % i for i in range(50)]
    tt = torch.tensor([1.0 if i % 2 else math.nan for i in range(50)], dtype=torch.float64, device="cuda")
    dkv.set_batch(md5_words(keys), torch.arange(50, dtype=torch.int64, device="cuda"), 0.0, tt)
    out, _ = dkv.contains_batch(md5_words(keys + ["zz"]), 0.5)
    assert out.tolist() == [1] * 50 + [0]
    out, val = dkv.contains_batch(md5_words(keys[:10]), 5.0)       # odd keys expired: deleted, no counters touched
    assert out.tolist() == [2 if i % 2 else 1 for i in range(10)]
    assert [v for o, v in zip(out.tolist(), val.tolist()) if o == 2] == [1, 3, 5, 7, 9]
    st = dkv.stats()
    assert (st["hits"], st["misses"], st["size"]) == (0, 0, 45)
    deleted, val = dkv.delete_batch(md5_words(["c0", "c0", "c1", "nope"]))
    assert sorted(deleted.tolist()[:2]) == [0, 1] and deleted.tolist()[2:] == [0, 0]
    n, purged = dkv.len_sync(5.0)
    assert n == 24 and sorted(purged) == list(range(11, 50, 2))
    dkv.clear()
    assert dkv.stats()["size"] == 0 and dkv.len_sync(0.0)[0] == 0
    hit, _ = dkv.get_batch(md5_words(keys[:4]), 0.0)
    assert hit.tolist() == [0, 0, 0, 0]


def test_c4_full_size_probe_and_fill_by_construction():
    """BASELINE.json configs[3]: 10M-entry table (LRU, default_ttl 300 s), probe batch 65536 at 90 % resident keys,
    then set the misses.  Too big for the Python oracle, so the expected result is built into the inputs:
    entries are inserted in id order (so LRU order == id order), the oldest 1 % are created 301 s before the
    probe (expired at probe time, strict '>'), the rest 1 s before."""
    from die import ops
    from die.kvstore import DeviceKV
    import bench
    N, B = 10_000_000, 65536
    kv = DeviceKV(N, 300.0)

    def dig(ids):
        kb = torch.from_numpy(bench.key_bytes_of(ids).reshape(-1)).cuda()
        off = torch.arange(ids.size + 1, dtype=torch.int64, device="cuda") * 15
        return ops.route_md5(ops.KeyBatch(kb, off), 1)[1]

    n_old = N // 100
    chunk = 1 << 20
    for a in range(0, N, chunk):
        m = min(chunk, N - a)
        ids = np.arange(a, a + m, dtype=np.int64)
        # ids < n_old are created at t=0, the others at t=300 (chunks never straddle: n_old < chunk handled by split)
        if a < n_old < a + m:
            cut = n_old - a
            kv.set_batch(dig(ids[:cut]), torch.from_numpy(ids[:cut]).cuda(), 0.0)
            kv.set_batch(dig(ids[cut:]), torch.from_numpy(ids[cut:]).cuda(), 300.0)
        else:
            kv.set_batch(dig(ids), torch.from_numpy(ids).cuda(), 0.0 if a + m <= n_old else 300.0)
    assert kv.sync_status()[0] == 0
    st = kv.stats()
    assert (st["size"], st["evictions"]) == (N, 0)

    rng = np.random.default_rng(4)
    resident = rng.integers(0, N, int(B * 0.9))
    absent = rng.integers(N, 2 * N, B - resident.size)
    ids = np.concatenate([resident, absent])
    rng.shuffle(ids)
    now = 301.0                                   # age of the old 1 %: 301 s > ttl 300 s -> expired; others: 1 s
    hit, value = kv.get_batch(dig(ids), now)
    hit, value = hit.cpu().numpy(), value.cpu().numpy()
    want_hit = (ids < N) & (ids >= n_old)
    assert np.array_equal(hit == 1, want_hit)
    assert np.array_equal(value[want_hit], ids[want_hit])                       # the stored handle is the id
    expired_ids = np.unique(ids[(ids < n_old)])
    assert int((hit == 2).sum()) == expired_ids.size                          # each expired entry deleted exactly once
    st = kv.stats()
    assert st["hits"] == int(want_hit.sum()) and st["misses"] == B - int(want_hit.sum())
    assert st["size"] == N - expired_ids.size

    # fill: set every miss (request order, duplicates included) -> the table is full again after expired.size inserts,
    # every further new key evicts the LRU head: untouched ids in increasing order (hits moved to MRU, expired gone)
    miss_ids = ids[~want_hit]
    kv.set_batch(dig(miss_ids), torch.from_numpy(miss_ids).cuda(), now)
    status, _, slow = kv.sync_status()
    assert status == 0 and slow == 0
    n_new = np.unique(miss_ids).size
    want_evictions = n_new - expired_ids.size
    st = kv.stats()
    assert st["size"] == N and st["evictions"] == want_evictions
    touched = np.zeros(N, dtype=bool)
    touched[ids[want_hit]] = True
    touched[:n_old][np.isin(np.arange(n_old), expired_ids)] = True             # deleted, not evictable
    untouched_order = np.nonzero(~touched)[0]                                   # LRU order of the evictable entries
    victims, survivors = untouched_order[:want_evictions], untouched_order[want_evictions:want_evictions + 4096]
    gone, _ = kv.contains_batch(dig(victims[-4096:]), now)
    kept, _ = kv.contains_batch(dig(survivors), now)
    assert not gone.any() and bool((kept.cpu().numpy()[survivors >= n_old] == 1).all())


@pytest.mark.parametrize("policy,max_size,space,n_batches,max_b", [("fifo", 500, 900, 40, 300), ("fifo", 37, 60, 50, 30),
                                                                   ("lfu", 64, 120, 40, 40), ("lfu", 9, 20, 40, 12)])
def test_random_batches_match_oracle_lfu_fifo(policy, max_size, space, n_batches, max_b):
    """The batched device API (duplicate keys inside a batch included) against the oracle applied op by op, for the
    FIFO and LFU policies: hit / miss and values per op, counters, and the eviction order at the end."""
    from die.kvstore import DeviceKV
    rng = np.random.default_rng(max_size + space)
    dkv = DeviceKV(max_size, None, policy=policy)
    clock = [0.0]
    port = rp.PortKVCache(max_size, None, clock=lambda: clock[0], policy=policy)
    now, next_val = 10.0, 1
    for b in range(n_batches):
        now += 1.0
        clock[0] = now
        B = int(rng.integers(1, max_b + 1))
        keys = ["k%d" % i for i in rng.integers(0, space, B)]
        dig = md5_words(keys)
        if rng.random() < 0.5:
            hit, value = dkv.get_batch(dig, now)
            want = [port.get(k, None) for k in keys]
            assert [v if h == 1 else None for h, v in zip(hit.tolist(), value.tolist())] == want, b
        else:
            values = list(range(next_val, next_val + B))
            next_val += B
            dkv.set_batch(dig, torch.tensor(values, dtype=torch.int64, device="cuda"), now)
            for k, v in zip(keys, values):
                port.set(k, v)
            assert dkv.sync_status()[0] == 0
        st, ps = dkv.stats(), port.stats()
        assert (st["hits"], st["misses"], st["evictions"], st["size"]) == (ps["hits"], ps["misses"], ps["evictions"], ps["size"]), b
    d = dkv.dump()
    order = [bytes(x) for x in d["digests"].numpy().view(np.uint8).reshape(-1, 16)]
    if policy == "fifo":
        assert order == [hashlib.md5(k.encode()).digest() for k in port.cache]
    else:
        assert order == [hashlib.md5(k.encode()).digest() for k in port.counts]
        assert d["lfu_count"].tolist() == list(port.counts.values())
    assert port.stats()["evictions"] > 0
