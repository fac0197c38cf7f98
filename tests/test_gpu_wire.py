"""GPU: binary request frames -> device CSR buffers (die_frame_decode, csrc/wire.cu) against an independent CPU
reading of the same bytes, including malformed frames, and end to end into the dispatch pipeline."""
import json
import struct

import numpy as np
import pytest
import torch

from oracle import wire as owire

pytestmark = pytest.mark.gpu


def check(buf, dim):
    from die import wire
    got = wire.decode_frames(buf, dim)
    offs = wire.split_frames(buf)
    want = owire.decode(buf, offs, dim)
    n = len(want)
    assert len(got.keys) == n
    ko = got.keys.key_off.cpu().numpy()
    kb = got.keys.key_bytes.cpu().numpy().tobytes()
    assert got.status.cpu().tolist() == [w[0] for w in want]
    assert got.n_bad == sum(1 for w in want if w[0] != 0)
    assert [int(v) & 0xFFFFFFFFFFFFFFFF for v in got.request_id.cpu().tolist()] == [w[1] for w in want]
    assert got.queue.cpu().tolist() == [w[2] for w in want]
    assert [kb[ko[i]:ko[i + 1]] for i in range(n)] == [w[3] for w in want]
    assert np.array_equal(got.payload.cpu().numpy(), np.stack([w[4] for w in want]) if n else np.zeros((0, dim), np.float32))
    return got


def test_frames_decode_bit_exact():
    from die import wire
    rng = np.random.default_rng(3)
    keys = ["", "a", "user-0000000042", "k" * 55, "k" * 56, "k" * 63, "k" * 64, "k" * 65, "k" * 119, "k" * 200, "ключ-é-鍵"]
    frames = []
    for i in range(3000):
        key = keys[i % len(keys)] if i % 3 else "user-%010d" % rng.integers(0, 10**7)
        nf = int(rng.integers(0, 129))
        frames.append(wire.encode_request(int(rng.integers(0, 2**63)) * 2 + 1, key, int(rng.integers(0, 5)),
                                          rng.standard_normal(nf).astype(np.float32)))
    got = check(b"".join(frames), 128)
    assert got.n_bad == 0
    check(b"", 8)                                  # nothing received
    check(wire.encode_request(7, "x", 0, []), 4)   # a single frame without inputs


def test_malformed_frames_are_flagged_not_trusted():
    from die import wire
    good = [wire.encode_request(i + 1, "key-%d" % i, 1, np.arange(8, dtype=np.float32) + i) for i in range(40)]
    frames = list(good)
    bad_magic = bytearray(good[3]); bad_magic[0] ^= 0xFF
    frames[3] = bytes(bad_magic)
    lying_key = bytearray(good[7]); struct.pack_into("<H", lying_key, 16, 60000)       # key_len beyond the frame
    frames[7] = bytes(lying_key)
    lying_n = bytearray(good[11]); struct.pack_into("<I", lying_n, 20, 1 << 30)         # n_floats beyond the frame
    frames[11] = bytes(lying_n)
    frames[20] = wire.encode_request(99, "wide", 0, np.ones(9, np.float32))             # more inputs than the row holds
    got = check(b"".join(frames), 8)
    st = got.status.cpu().tolist()
    assert [i for i, s in enumerate(st) if s] == [3, 7, 11, 20] and st[20] == wire.FRAME_TOO_WIDE
    # a length prefix that leaves the buffer: the receive loop stops there, the tail is one bad frame
    cut = bytearray(b"".join(good[:5])); struct.pack_into("<I", cut, len(good[0]) * 2 + 4, 1 << 20)
    got = check(bytes(cut), 8)
    assert got.status.cpu().tolist() == [0, 0, 1]


def test_frames_stand_for_the_reference_json_request_and_feed_the_pipeline():
    """The JSON the reference's worker parses (worker.py:96-100) and the frame carry the same fields; decoded frames go
    straight into the dispatch pipeline and give what the string / array path gives."""
    from die import ops, wire
    from die.pipeline import DispatchPipeline
    rng = np.random.default_rng(9)
    B, D = 512, 32
    reqs = [{"model": "m", "version": "1", "inputs": rng.standard_normal(D).astype(np.float32).tolist(),
             "request_id": "%d" % (1000 + i), "key": "user-%010d" % rng.integers(0, 300)} for i in range(B)]
    as_json = [json.loads(json.dumps(r)) for r in reqs]   # what the reference receives
    buf = b"".join(wire.encode_request(int(r["request_id"]), r["key"], 0, r["inputs"]) for r in as_json)
    got = wire.decode_frames(buf, D)
    assert got.n_bad == 0 and got.request_id.cpu().tolist() == [int(r["request_id"]) for r in reqs]
    a = DispatchPipeline(400, B, D, model="echo", default_ttl=60.0)
    b = DispatchPipeline(400, B, D, model="echo", default_ttl=60.0)
    x = torch.from_numpy(np.asarray([r["inputs"] for r in reqs], dtype=np.float32)).cuda()
    for now in (1.0, 2.0):
        o1, h1 = a.run_device(got.keys, got.payload, now)
        o2, h2 = b.run_device(ops.KeyBatch.from_strings([r["key"] for r in reqs]), x, now)
        assert torch.equal(o1, o2) and torch.equal(h1, h2)
    assert a.kv.stats() == b.kv.stats() and a.kv.stats()["hits"] > 0
