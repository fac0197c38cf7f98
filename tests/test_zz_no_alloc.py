"""die_kv_set_batch_noalloc: the batched cache fill for hosts that may not allocate device memory while
serving (all HBM reserved at start-up, or a section that cannot afford cudaMalloc's device-wide sync).

The entry point keeps its per-op scratch in a fixed array inside the library instead of a cudaMalloc'd
buffer.  This test drives it through the C-ABI twice in a row and requires exactly what the regular
die_kv_set_batch produces on a twin table: same dropped values, same LRU order, same entry
metadata and counters (bit-exact, like every other cache test here), and what the CPU restatement of the
reference's KVCache.set holds after the same ops.  The second call is smaller than the first, so it runs over
whatever the first call left in the shared scratch.  Collected last on purpose."""
import ctypes
import hashlib
import math

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def md5_words(keys):
    raw = b"".join(hashlib.md5(k.encode()).digest() for k in keys)
    return torch.from_numpy(np.frombuffer(raw, dtype=np.int32).reshape(-1, 4).copy()).cuda()


def call_set(lib, name, kv, dig, values, ttl, now):
    from die._cabi import ptr, stream_ptr
    M = dig.shape[0]
    slot = torch.empty(M, dtype=torch.int32, device="cuda")
    dropped = torch.full((2 * M,), -1, dtype=torch.int64, device="cuda")
    rc = getattr(lib, name)(kv._h, ptr(dig), ptr(values), ptr(ttl), M, now, 0, ptr(slot), ptr(dropped), stream_ptr())
    return rc, slot, dropped


def state_of(kv):
    d = kv.dump()
    return {k: v.tolist() for k, v in d.items()}


def test_noalloc_set_matches_regular_set_on_two_calls_in_a_row():
    from die import _cabi
    from die.kvstore import DeviceKV
    lib = _cabi.require_cuda()
    from oracle import ref_port as rp
    ref, sut = DeviceKV(5000, 30.0), DeviceKV(5000, 30.0)   # per-launch limit of such a table: 2048 ops
    clock = [0.0]
    port = rp.PortKVCache(5000, 30.0, clock=lambda: clock[0])  # kvstore.py:104-134 applied op by op
    rng = np.random.default_rng(42)
    now, next_val = 10.0, 1
    # back-to-back calls: fill (duplicates inside each batch), overflow the 5000 entries (evictions),
    # then a much smaller batch of overwrites + new keys that runs over the previous call's scratch
    for n_ops, space in ((2000, 2500), (2000, 5000), (2000, 9000), (250, 9000)):
        now += 1.5
        keys = ["na%d" % i for i in rng.integers(0, space, n_ops)]
        dig = md5_words(keys)
        values = torch.arange(next_val, next_val + n_ops, dtype=torch.int64, device="cuda")
        next_val += n_ops
        ttl = torch.tensor([math.nan if i % 3 else 5.0 for i in range(n_ops)], dtype=torch.float64, device="cuda")
        rc_a, slot_a, drop_a = call_set(lib, "die_kv_set_batch", ref, dig, values, ttl, now)
        rc_b, slot_b, drop_b = call_set(lib, "die_kv_set_batch_noalloc", sut, dig, values, ttl, now)
        assert rc_a == 0 and rc_b == 0, lib.die_last_error()
        st_a, nd_a, _ = ref.sync_status()
        st_b, nd_b, _ = sut.sync_status()
        assert (st_a, nd_a) == (st_b, nd_b) and st_b == 0
        # which free slot of a probe chain an insert wins depends on thread timing, so slot numbers may differ
        # between the twins; what must hold is one slot per distinct key, shared by its duplicates
        for slot in (slot_a.tolist(), slot_b.tolist()):
            by_key = {}
            for k, s_ in zip(keys, slot):
                assert s_ >= 0 and by_key.setdefault(k, s_) == s_
            assert len(set(by_key.values())) == len(by_key)
        assert sorted(drop_a[:nd_a].tolist()) == sorted(drop_b[:nd_b].tolist())
        assert ref.stats() == sut.stats()
        assert state_of(ref) == state_of(sut)
        # ... and against the oracle directly, not only against the twin: LRU order, values, counters
        clock[0] = now
        for k, v, t in zip(keys, values.tolist(), ttl.tolist()):
            port.set(k, v, ttl=None if math.isnan(t) else t)
        d = sut.dump()
        order = [bytes(x) for x in d["digests"].numpy().view(np.uint8).reshape(-1, 16)]
        assert order == [hashlib.md5(k.encode()).digest() for k in port.lru_order()]
        assert d["values"].tolist() == [port.cache[k].value for k in port.lru_order()]
        ps, ss = port.stats(), sut.stats()
        assert (ss["size"], ss["evictions"]) == (ps["size"], ps["evictions"])
        # and the filled table answers like its twin
        probe_keys = ["na%d" % i for i in rng.integers(0, space, 300)]
        probe = md5_words(probe_keys)
        now += 0.5
        hit_a, val_a = ref.get_batch(probe, now)
        hit_b, val_b = sut.get_batch(probe, now)
        assert torch.equal(hit_a, hit_b) and torch.equal(val_a, val_b)
        clock[0] = now
        want = [port.get(k, None) for k in probe_keys]   # the gets move their hits to the MRU end in the oracle too
        assert [v if h == 1 else None for h, v in zip(hit_b.tolist(), val_b.tolist())] == want


def test_noalloc_set_rejects_batches_beyond_its_fixed_scratch():
    from die import _cabi
    from die.kvstore import DeviceKV
    lib = _cabi.require_cuda()
    kv = DeviceKV(1 << 20, None)          # large table: its own per-launch limit is above the scratch limit
    M = 65536 + 1
    dig = torch.zeros((M, 4), dtype=torch.int32, device="cuda")
    rc, _, _ = call_set(lib, "die_kv_set_batch_noalloc", kv, dig, None, None, 1.0)
    assert rc == _cabi.DIE_ERR_TOO_LARGE
    assert b"65536" in lib.die_last_error()
    assert kv.stats()["size"] == 0        # nothing was applied
    ok = md5_words(["fits-%d" % i for i in range(65536)])
    rc, slot, _ = call_set(lib, "die_kv_set_batch_noalloc", kv, ok, None, None, 2.0)
    assert rc == 0, lib.die_last_error()
    assert kv.sync_status()[0] == 0 and kv.stats()["size"] == 65536 and int(slot.min()) >= 0


def test_noalloc_sets_interleaved_with_replayed_pipeline_steps_keep_the_host_bounds_sound():
    """A pipeline-owned table (die_pipeline_kv) refreshes its free-slot bound from a device snapshot that counts the
    inserts enqueued since (`ops_enq`).  No-alloc sets on that table must be part of the count: interleave them with
    graph-replayed queues of batches and require that housekeeping still keeps the table under its load limit."""
    from die import _cabi
    from die.pipeline import DispatchPipeline
    lib = _cabi.require_cuda()
    B, D, kv_max = 256, 32, 1 << 16   # table of 2^18 slots: 16-step replays fit its reserve
    rng = np.random.default_rng(5)
    pipe = DispatchPipeline(kv_max, B, D, model="echo", default_ttl=None)
    off = torch.arange(B + 1, dtype=torch.int64, device="cuda") * 15
    out = torch.empty((4, B, D), dtype=torch.float32, device="cuda")
    hit = torch.empty((4, B), dtype=torch.uint8, device="cuda")
    now = 1.0
    for rnd in range(40):
        ids = rng.integers(0, 10 ** 9, (32, B))
        kb = np.frombuffer("".join("user-%010d" % i for i in ids.reshape(-1)).encode(), dtype=np.uint8).reshape(32, B * 15)
        keys = torch.from_numpy(np.concatenate([kb, np.zeros((32, 16), np.uint8)], axis=1)).cuda()
        x = torch.from_numpy(rng.standard_normal((32, B, D), dtype=np.float32)).cuda()
        pipe.run_device_many(keys, off, x, now, 0.01, out, hit)
        now += 1.0
        for _ in range(4):
            # no values passed: the pipeline's table hands out its own result rows, B spare rows per batch
            dig = md5_words(["na-%d-%d" % (rnd, i) for i in rng.integers(0, 10 ** 9, B)])
            rc, _, _ = call_set(lib, "die_kv_set_batch_noalloc", pipe.kv, dig, None, None, now)
            if rc == _cabi.DIE_ERR_NOMEM:         # housekeeping due: run it where allocation is allowed, retry
                pipe.kv.maintain(force=False)
                rc, _, _ = call_set(lib, "die_kv_set_batch_noalloc", pipe.kv, dig, None, None, now)
            assert rc == 0, lib.die_last_error()
        st = pipe.kv.stats()
        assert st["size"] + st["tombstones"] <= st["capacity"] // 2, (rnd, st)  # the load limit the bounds protect
    assert pipe.kv.sync_status()[0] == 0 and pipe.kv.stats()["size"] == kv_max
