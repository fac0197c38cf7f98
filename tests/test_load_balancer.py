"""CPU: die.load_balancer.LoadBalancer against sessions recorded from the reference's own class
(tests/golden/make_golden.py::gen_load_balancer): every pick of every strategy, pinned requests, failures that take a
worker out, unregister / re-register, statistics."""
import asyncio
import os
import random
import sys

import pytest

from conftest import load_golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [os.path.join(ROOT, "distributed-inference-engine_b200")]


@pytest.mark.parametrize("strategy", ["round_robin", "least_connections", "random", "least_latency"])
def test_load_balancer_session_matches_reference(strategy):
    from die.load_balancer import LoadBalancer, LoadBalancerStrategy
    case = load_golden("load_balancer.json")[strategy]

    async def session():
        random.seed(4242)
        lb = LoadBalancer(strategy=LoadBalancerStrategy(strategy), max_failures=2)
        log = [await lb.get_worker()]
        for i in range(4):
            lb.register_worker("w%d" % i, "127.0.0.1:%d" % (9100 + i))
        for op in case["script"]:
            if op[0] == "get":
                got = await lb.get_worker(op[1])
                log.append(list(got) if got else None)
            elif op[0] == "stats":
                await lb.update_stats(op[1], op[2], op[3])
                log.append("ok")
            elif op[0] == "conn":
                if op[1] in lb.worker_stats:
                    lb.worker_stats[op[1]].active_connections = op[2]
                log.append("ok")
            elif op[0] == "unregister":
                log.append(lb.unregister_worker(op[1]))
            else:
                lb.register_worker(op[1], "127.0.0.1:9999")
                log.append("ok")
        stats = {w: {k: v for k, v in s.items() if k != "last_seen"} for w, s in lb.get_all_stats().items()}
        return log, stats, lb.get_worker_stats("nope"), repr(lb)

    log, stats, unknown, rep = asyncio.run(session())
    assert log == case["log"]
    assert stats == case["stats"]
    assert unknown == case["unknown"] and rep == case["repr"]


def test_health_check_marks_and_recovers_workers():
    """The background probe (load_balancer.py:293-345) with a scripted probe instead of a socket: consecutive failures
    take a worker out of rotation, one success brings it back; stop() cancels the loop."""
    from die.load_balancer import LoadBalancer
    down = {"b:2"}

    async def probe(address, timeout):
        if address in down:
            raise ConnectionError("refused")

    async def run():
        lb = LoadBalancer(health_check_interval=0.01, max_failures=2, probe=probe)
        lb.register_worker("a", "a:1")
        lb.register_worker("b", "b:2")
        await lb.start()
        await asyncio.sleep(0.08)
        picks = {(await lb.get_worker())[0] for _ in range(6)}
        pinned = await lb.get_worker("b")
        healthy_b = lb.get_worker_stats("b")["healthy"]
        down.clear()
        await asyncio.sleep(0.05)
        back = lb.get_worker_stats("b")["healthy"]
        await lb.stop()
        return picks, pinned, healthy_b, back, lb.get_worker_stats("a")["request_count"]

    picks, pinned, healthy_b, back, probes_a = asyncio.run(run())
    assert picks == {"a"} and pinned is None and healthy_b is False and back is True and probes_a >= 3


def test_wire_frames_encode_and_split_on_the_host():
    """die.wire's host half (no GPU): frames carry the fields of the reference's JSON request (worker.py:96-100), the
    receive-loop walk finds every frame start from the length prefixes and stops at a prefix that leaves the buffer;
    the CPU reading of the same bytes (oracle/wire.py, what the GPU decode is checked against) gets the fields back."""
    import struct

    import numpy as np
    from die import wire
    from oracle import wire as owire
    reqs = [(7, "user-0000000001", 0, [1.0, 2.5]), (2**64 - 1, "", 3, []), (9, "ключ-é", 1, list(range(9)))]
    frames = [wire.encode_request(*r) for r in reqs]
    assert all(len(f) % 4 == 0 for f in frames)
    buf = b"".join(frames)
    offs = wire.split_frames(buf)
    assert offs.tolist() == [0, len(frames[0]), len(frames[0]) + len(frames[1]), len(buf)]
    got = owire.decode(buf, offs, 16)
    for (rid, key, q, x), (st, grid, gq, gkey, row) in zip(reqs, got):
        assert (st, grid, gq, gkey) == (0, rid, q, key.encode())
        assert row[:len(x)].tolist() == [float(v) for v in x] and not row[len(x):].any()
    # a prefix that points past the buffer ends the walk: one trailing bad frame
    cut = bytearray(buf)
    struct.pack_into("<I", cut, len(frames[0]) + 4, 1 << 20)
    offs = wire.split_frames(bytes(cut))
    assert offs.tolist() == [0, len(frames[0]), len(buf)]
    assert [g[0] for g in owire.decode(bytes(cut), offs, 16)] == [0, 1]
    assert owire.decode(wire.encode_request(1, "k", 0, np.ones(17)), [0, 24 + 4 + 68], 16)[0][0] == 2   # too wide
    with pytest.raises(ValueError):
        wire.encode_request(1, "k" * 70000, 0, [])
