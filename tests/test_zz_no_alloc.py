"""die_kv_set_batch_noalloc: the batched cache fill for hosts that may not allocate device memory while
serving (all HBM reserved at start-up, or a section that cannot afford cudaMalloc's device-wide sync).

The entry point keeps its per-op scratch in a fixed array inside the library instead of a cudaMalloc'd
buffer.  This test drives it through the C-ABI twice in a row and requires exactly what the regular
die_kv_set_batch produces on a twin table: same dropped values, same LRU order, same entry
metadata and counters (bit-exact, like every other cache test here).  The second call is smaller than the
first, so it runs over whatever the first call left in the shared scratch.  Collected last on purpose."""
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
    ref, sut = DeviceKV(5000, 30.0), DeviceKV(5000, 30.0)   # per-launch limit of such a table: 2048 ops
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
        # and the filled table answers like its twin
        probe = md5_words(["na%d" % i for i in rng.integers(0, space, 300)])
        now += 0.5
        hit_a, val_a = ref.get_batch(probe, now)
        hit_b, val_b = sut.get_batch(probe, now)
        assert torch.equal(hit_a, hit_b) and torch.equal(val_a, val_b)


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
