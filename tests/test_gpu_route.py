"""GPU parity: K1 route (MD5 mod N + health gate + failover) through the C-ABI vs the reference goldens
and the C oracle.  Bit-exact."""
import numpy as np
import pytest
import torch

from conftest import load_golden
from oracle import c_oracle
from oracle import ref_port as rp

pytestmark = pytest.mark.gpu


def dev_route(keys, n, routable=None, healthy=None, failover=True, counts=False):
    from die import ops
    kb = ops.KeyBatch.from_strings(keys)
    r = None if routable is None else torch.tensor(routable, dtype=torch.uint8, device="cuda")
    h = None if not healthy else torch.tensor(healthy, dtype=torch.int32, device="cuda")
    shard, dig, cnt = ops.route_md5(kb, n, r, h, failover, want_digests=True, want_counts=counts)
    torch.cuda.synchronize()
    return shard.cpu().numpy(), dig.cpu().numpy().view(np.uint8).reshape(-1, 16), None if cnt is None else cnt.cpu().numpy()


def test_golden_digests_and_shard_indices():
    g = load_golden("route.json")
    keys = g["keys"]
    for n, want in g["shard_index"].items():
        shard, dig, cnt = dev_route(keys, int(n), counts=True)
        assert [d.tobytes().hex() for d in dig] == g["md5"]
        assert shard.tolist() == want, n
        assert cnt.tolist() == np.bincount(np.array(want), minlength=int(n)).tolist()


def test_registry_and_router_drop_in_against_golden():
    from die import ModelRegistry, Router
    g = load_golden("route.json")
    reg = ModelRegistry()
    reg.register_model("m", "1", "/p", {}, {})
    for sid in g["out_of_order"]["shard_ids"]:
        reg.add_shard("m", "1", sid, f"w{sid}")
    got = [s.shard_id for s in reg.get_shards_for_keys("m", "1", g["keys"])]
    assert got == g["out_of_order"]["routed_ids"]  # list position, not shard_id
    assert reg.get_shard_for_key("m", "1", "user-123").shard_id == got[0]
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
        routed = router.route("m", "1", g["keys"][:48], prefer_healthy=case["prefer_healthy"]).tolist()
        assert routed == case["routed"], case
        one = router.route_request("m", "1", g["keys"][0], prefer_healthy=case["prefer_healthy"])
        assert (-1 if one is None else one.shard_id) == case["routed"][0]
    assert router.route_request("nope", "1", "k") is None


def test_random_keys_all_lengths_vs_c_oracle():
    rng = np.random.default_rng(1)
    alphabet = np.array(list("abcdefghijklmnopqrstuvwxyz0123456789-_é中"))
    keys = ["".join(rng.choice(alphabet, size=int(n))) for n in rng.integers(0, 140, 20000)]
    keys += ["", "", "x" * 55, "x" * 56, "x" * 119, "x" * 120] * 3
    kb, off = c_oracle.encode(keys)
    for n, routable, healthy, failover in [(3, None, None, True), (8, [1, 0, 1, 1, 0, 1, 1, 1], [0, 2, 3, 5], True),
                                           (8, [0] * 8, [], True), (5, [1, 0, 0, 1, 1], [3], False), (1000, None, None, True)]:
        want_shard, want_dig = c_oracle.route(kb, off, n, routable, healthy, failover)
        shard, dig, _ = dev_route(keys, n, routable, healthy, failover)
        assert np.array_equal(dig, want_dig)
        assert np.array_equal(shard, want_shard), n


def test_long_keys_take_the_global_path():
    keys = [("k%05d" % i) * 60 for i in range(700)]  # 360 bytes each: a CTA's span exceeds the smem stage
    kb, off = c_oracle.encode(keys)
    want_shard, want_dig = c_oracle.route(kb, off, 7)
    shard, dig, _ = dev_route(keys, 7)
    assert np.array_equal(dig, want_dig) and np.array_equal(shard, want_shard)


def test_full_size_batch_1m_keys():
    ids = np.random.default_rng(0).integers(0, 10 ** 7, 1 << 20)
    keys = ["user-%010d" % i for i in ids]
    kb, off = c_oracle.encode(keys)
    want_shard, want_dig = c_oracle.route(kb, off, 8)
    shard, dig, cnt = dev_route(keys, 8, counts=True)
    assert np.array_equal(dig, want_dig) and np.array_equal(shard, want_shard)
    assert cnt.tolist() == np.bincount(want_shard, minlength=8).tolist()


def test_empty_batch_and_zero_shards():
    shard, dig, _ = dev_route([], 4)
    assert shard.size == 0 and dig.size == 0
    shard, _, _ = dev_route(["a", "b"], 0)
    assert shard.tolist() == [-1, -1]
    shard, _, _ = dev_route(["", "", ""], 2)
    assert shard.tolist() == [rp.shard_index("", 2)] * 3
