"""Binary request frames (include/die.h, die_frame_header_t): the length-prefixed replacement of the JSON request the
reference's worker reads from its socket (src/worker.py:92-100), decoded on the GPU into the path's CSR buffers.

    frame = encode_request(request_id, key, queue, inputs)        # client side
    batch = decode_frames(buf)                                    # server side: bytes of any number of frames

The transport itself stays with the caller (a socket reader knows where frames start from `frame_bytes`)."""
from __future__ import annotations

import struct
from dataclasses import dataclass
from typing import Sequence

import numpy as np
import torch

from . import _cabi, ops

MAGIC = 0x31454944
HEADER = struct.Struct("<IIQHHI")   # magic, frame_bytes, request_id, key_len, queue, n_floats
FRAME_OK, FRAME_BAD, FRAME_TOO_WIDE = 0, 1, 2


def encode_request(request_id: int, key: str, queue: int, inputs: Sequence[float]) -> bytes:
    kb = key.encode("utf-8")
    if len(kb) > 0xFFFF:
        raise ValueError("key longer than 65535 bytes")
    x = np.asarray(inputs, dtype="<f4")
    pad = (-len(kb)) % 4
    n = HEADER.size + len(kb) + pad + 4 * x.size
    return HEADER.pack(MAGIC, n, request_id, len(kb), queue, x.size) + kb + b"\0" * pad + x.tobytes()


def split_frames(buf: bytes) -> np.ndarray:
    """Frame starts of a buffer of back-to-back frames, from the length prefixes (what a receive loop does).  A prefix
    that would leave the buffer, or is not a positive multiple of 4, ends the walk: the rest is one (bad) frame."""
    offs, o, n = [0], 0, len(buf)
    while o < n:
        size = struct.unpack_from("<I", buf, o + 4)[0] if o + 8 <= n else 0
        if size < HEADER.size or size % 4 or o + size > n:
            o = n
        else:
            o += size
        offs.append(o)
    return np.asarray(offs, dtype=np.int64)


@dataclass
class DecodedBatch:
    keys: ops.KeyBatch
    payload: torch.Tensor      # [n, dim] fp32, zero-padded
    request_id: torch.Tensor   # [n] int64 (the frames' u64 ids)
    queue: torch.Tensor        # [n] int32, -1 for frames that are not OK
    status: torch.Tensor       # [n] uint8
    n_bad: int


def decode_frames(buf: bytes, dim: int) -> DecodedBatch:
    """One H2D copy of the received bytes, two kernels (die_frame_decode)."""
    lib = _cabi.require_cuda()
    off = split_frames(buf)
    n = len(off) - 1
    raw = np.frombuffer(buf + b"\0" * ((-len(buf)) % 4 + 4), dtype=np.uint8)
    d_buf = torch.from_numpy(raw.copy()).cuda()
    d_off = torch.from_numpy(off).cuda()
    key_bytes = torch.zeros(len(buf) + 16, dtype=torch.uint8, device="cuda")   # keys are a subset of the frames' bytes
    key_off = torch.empty(n + 1, dtype=torch.int64, device="cuda")
    payload = torch.empty((n, dim), dtype=torch.float32, device="cuda")
    rid = torch.empty(n, dtype=torch.int64, device="cuda")
    queue = torch.empty(n, dtype=torch.int32, device="cuda")
    status = torch.empty(n, dtype=torch.uint8, device="cuda")
    n_bad = torch.zeros(1, dtype=torch.int32, device="cuda")
    ptr = _cabi.ptr
    _cabi.check(lib.die_frame_decode(ptr(d_buf), ptr(d_off), n, dim, ptr(key_bytes), ptr(key_off), ptr(payload), ptr(rid),
                                     ptr(queue), ptr(status), ptr(n_bad), _cabi.stream_ptr()))
    return DecodedBatch(ops.KeyBatch(key_bytes, key_off), payload, rid, queue, status, int(n_bad.item()))
