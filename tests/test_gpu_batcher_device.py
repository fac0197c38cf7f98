"""GPU: the device-resident Batcher (die_batcher_*, csrc/batcher.cu) against the reference's flush schedule -- the
golden ticks recorded from the reference's own Batcher on virtual time, and seeded differential runs against the
CPU restatement (oracle.ref_port.PortTickBatcher) with the mock linear model per key."""
import numpy as np
import pytest
import torch

from conftest import load_golden, replay_ticks
from oracle import models
from oracle import ref_port as rp

pytestmark = pytest.mark.gpu
TICK_CASES = ["one_key_small", "two_keys", "five_keys_bursts", "mostly_latency", "mostly_size"]


def dev(a, dtype):
    return torch.from_numpy(np.asarray(a, dtype=dtype)).cuda()


@pytest.mark.parametrize("name", TICK_CASES)
def test_device_batcher_replays_reference_ticks(name):
    """Echo model (FakeModel): every batch the reference flushed, with the same members in the same order, at the same
    point of the tick sequence; every request gets its own payload back; the counters of get_stats()."""
    from die.device_batcher import DeviceBatcher
    case = load_golden("tick_batcher.json")[name]
    keys = [tuple(k.split(":")) for k in case["keys"]]
    qof = {k: i for i, k in enumerate(case["keys"])}
    n_req = sum(len(t["arrivals"]) for t in case["ticks"])
    D = 8
    b = DeviceBatcher(keys, case["max_batch_size"], case["max_latency_ms"], dim_in=D, max_arrivals=256)
    out = torch.full((n_req, D), -1.0, device="cuda")
    batch_of = torch.full((n_req,), -1, dtype=torch.int32, device="cuda")
    payload_of = np.arange(n_req * D, dtype=np.float32).reshape(n_req, D)
    state = {}

    def tick(now, arrivals, force):
        if force:
            st = b.get_stats()
            state["before_stop"] = {k: st[k] for k in case["stats_before_stop"]}
        if arrivals:
            ids = [rid for _k, rid in arrivals]
            b.tick(now, dev([qof[k] for k, _r in arrivals], np.int32), dev(ids, np.int32), dev(payload_of[ids], np.float32),
                   out, batch_of, force)
        else:
            b.tick(now, None, None, None, out, batch_of, force)
        return [(f["key"], f["ids"], f["trigger"]) for f in b.last_flushes()]

    replay_ticks(case, tick)
    assert state["before_stop"] == case["stats_before_stop"]
    st = b.get_stats()
    assert {k: st[k] for k in case["stats"]} == case["stats"]
    assert np.array_equal(out.cpu().numpy(), payload_of)          # UNPACK: result i -> request i (batcher.py:225-227)
    got_batch = batch_of.cpu().numpy()
    for j, f in enumerate(case["flushed"]):
        assert np.all(got_batch[f["ids"]] == j), j


@pytest.mark.parametrize("seed,n_queues,max_batch,latency_ms,max_per_tick", [
    (1, 5, 64, 3.5, 1500), (2, 1, 4096, 20.5, 6000), (3, 19, 7, 1.5, 300), (4, 2, 1, 5.5, 40)])
def test_device_batcher_matches_port_with_linear_models(seed, n_queues, max_batch, latency_ms, max_per_tick):
    """Seeded traffic, one mock linear layer per key: flush list (key, members, trigger) per tick, outputs within the
    fp32 tolerance of oracle/models.py, counters; bursts that flush one key several times in a tick; idle ticks."""
    from die.device_batcher import DeviceBatcher
    rng = np.random.default_rng(seed)
    K, N = 64, 32
    keys = [("m%d" % q, "v1") for q in range(n_queues)]
    names = ["m%d:v1" % q for q in range(n_queues)]
    W = (rng.standard_normal((n_queues, N, K)) / np.sqrt(K)).astype(np.float32)
    bias = rng.standard_normal((n_queues, N)).astype(np.float32)
    b = DeviceBatcher(keys, max_batch, latency_ms, dim_in=K, dim_out=N, max_arrivals=max_per_tick, weights=W, bias=bias)
    port = rp.PortTickBatcher(max_batch, latency_ms)
    n_ticks = 40
    total = n_ticks * max_per_tick
    out = torch.zeros((total, N), device="cuda")
    batch_of = torch.full((total,), -1, dtype=torch.int32, device="cuda")
    X = rng.standard_normal((total, K)).astype(np.float32)
    qs_all = np.full(total, -1, dtype=np.int64)
    now, rid, n_batches = 0.0, 0, 0
    for t in range(n_ticks):
        now += float(rng.choice([0.001, 0.002, 0.004, 0.011])) + 0.0001
        n = 0 if rng.random() < 0.2 else int(rng.integers(1, max_per_tick + 1))
        force = t == n_ticks - 1
        qs = rng.integers(0, n_queues, n)
        ids = np.arange(rid, rid + n)
        qs_all[ids] = qs
        rid += n
        if n:
            b.tick(now, dev(qs, np.int32), dev(ids, np.int32), dev(X[ids], np.float32), out, batch_of, force)
        else:
            b.tick(now, None, None, None, out, batch_of, force)
        want = port.tick(now, [(names[q], int(i)) for q, i in zip(qs, ids)], force)
        got = b.last_flushes()
        assert [(f["key"], f["ids"], f["trigger"]) for f in got] == [(k, m, trig) for k, m, trig in want], t
        assert [f["batch"] for f in got] == list(range(n_batches, n_batches + len(want)))
        n_batches += len(want)
        st = b.get_stats()
        assert (st["total_requests"], st["total_batches"], st["total_batched_requests"]) == \
               (port.total_requests, port.total_batches, port.total_batched_requests)
        assert st["pending"] == {k: v for k, v in port.pending().items() if v}
    assert n_batches > 0 and rid > 0
    got = out.cpu().numpy()[:rid].astype(np.float64)
    for q in range(n_queues):
        sel = np.nonzero(qs_all[:rid] == q)[0]
        if len(sel) == 0:
            continue
        y = models.linear_f64(X[sel], W[q], bias[q])
        bound = models.linear_abs_bound(X[sel], W[q])
        assert np.all(np.abs(got[sel] - y) <= 1e-5 * bound + 1e-6), q
    assert np.all(batch_of.cpu().numpy()[:rid] >= 0)   # the forced last tick served everything


def test_device_batcher_argument_checks():
    from die import _cabi
    from die.device_batcher import DeviceBatcher
    with pytest.raises(ValueError, match="max_batch_size must be positive"):
        DeviceBatcher([("n", "1")], 0, 10.0)
    with pytest.raises(ValueError, match="max_latency_ms must be positive"):
        DeviceBatcher([("n", "1")], 4, 0.0)
    with pytest.raises(_cabi.DieError):
        DeviceBatcher([("n%d" % i, "1") for i in range(20)], 4, 1.0)      # more keys than the device batcher takes
    b = DeviceBatcher([("n", "1")], 4, 1.0, dim_in=8, max_arrivals=16)
    out = torch.zeros((64, 8), device="cuda")
    with pytest.raises(_cabi.DieError):
        b.tick(0.0, dev([0] * 17, np.int32), dev(range(17), np.int32), torch.zeros((17, 8), device="cuda"), out)
