"""Tensor-level wrappers over the C-ABI (one Python function per entry point of include/die.h).

torch is used for device memory and streams only; every byte of work is done by libdie's kernels.
"""
from __future__ import annotations

import math
from typing import List, Optional, Sequence

import numpy as np
import torch

from . import _cabi
from ._cabi import check, ptr, require_cuda, stream_ptr


_PINNED = {}


def pinned_staging(tag: str, nbytes: int) -> torch.Tensor:
    """A reusable pinned host buffer of at least nbytes (uint8), one per tag: allocating pinned memory costs a
    millisecond, so the Python-object edge of the drop-in classes stages its one H2D copy per batch through these."""
    buf = _PINNED.get(tag)
    if buf is None or buf.numel() < nbytes:
        cap = 1 << max(int(nbytes - 1).bit_length(), 12)
        buf = torch.empty(cap, dtype=torch.uint8).pin_memory()
        _PINNED[tag] = buf
    return buf


class KeyBatch:
    """CSR of UTF-8 key bytes on the device (`key.encode()` of model_registry.py:159)."""

    def __init__(self, key_bytes: torch.Tensor, key_off: torch.Tensor):
        self.key_bytes, self.key_off = key_bytes, key_off

    def __len__(self) -> int:
        return self.key_off.numel() - 1

    @staticmethod
    def encode_host(keys: Sequence[str]):
        """-> (uint8[total], int64[B+1]) numpy arrays: `key.encode()` of every key, back to back.

        ASCII keys (the common case) are encoded with one join + one encode, their byte lengths are their
        character counts; anything else falls back to one encode per key."""
        off = np.zeros(len(keys) + 1, dtype=np.int64)
        if not keys:
            return np.zeros(0, dtype=np.uint8), off
        joined = "".join(keys)
        raw = joined.encode()
        if len(raw) == len(joined):
            np.cumsum(np.fromiter(map(len, keys), dtype=np.int64, count=len(keys)), out=off[1:])
        else:
            enc = [k.encode() for k in keys]
            np.cumsum(np.fromiter(map(len, enc), dtype=np.int64, count=len(enc)), out=off[1:])
            raw = b"".join(enc)
        data = np.frombuffer(raw, dtype=np.uint8) if off[-1] else np.zeros(0, dtype=np.uint8)
        return data, off

    @classmethod
    def from_strings(cls, keys: Sequence[str], device="cuda") -> "KeyBatch":
        """Key strings -> device CSR with ONE host-to-device copy (bytes and offsets share a pinned staging buffer)."""
        require_cuda()
        data, off = cls.encode_host(keys)
        nb, no = int(data.size), off.size * 8
        pad = (-nb) % 8 + 16                      # 16 bytes of slack after the keys, offsets 8-byte aligned
        stage = pinned_staging("keys", nb + pad + no)
        host = stage.numpy()
        host[:nb] = data
        host[nb:nb + pad] = 0
        host[nb + pad:nb + pad + no] = off.view(np.uint8)
        dev = torch.empty(nb + pad + no, dtype=torch.uint8, device=device)
        dev.copy_(stage[:nb + pad + no], non_blocking=True)
        torch.cuda.current_stream().synchronize()  # the staging buffer is reused by the next call
        return cls(dev[:nb + pad], dev[nb + pad:].view(torch.int64))


def route_md5(keys: KeyBatch, n_shards: int, routable: Optional[torch.Tensor] = None,
              healthy_idx: Optional[torch.Tensor] = None, failover: bool = True, want_digests: bool = True,
              want_counts: bool = False):
    """-> (shard int32[B], digests uint32[B,4] | None, dest_counts int32[n_shards] | None)."""
    lib = require_cuda()
    B = len(keys)
    dev = keys.key_off.device
    shard = torch.empty(B, dtype=torch.int32, device=dev)
    digests = torch.empty((B, 4), dtype=torch.int32, device=dev) if want_digests else None
    counts = torch.empty(max(n_shards, 1), dtype=torch.int32, device=dev) if want_counts else None
    n_healthy = 0 if healthy_idx is None else healthy_idx.numel()
    check(lib.die_route_md5(ptr(keys.key_bytes), ptr(keys.key_off), B, n_shards, ptr(routable), ptr(healthy_idx),
                            n_healthy, 1 if failover else 0, ptr(shard), ptr(digests), ptr(counts), stream_ptr()))
    return shard, digests, (counts[:n_shards] if counts is not None else None)


def bucket_by_shard(shard: torch.Tensor, n_shards: int):
    """-> (offsets int64[n_shards+2], perm int32[B], pos int32[B]); stable, unroutable (-1) requests last."""
    lib = require_cuda()
    B = shard.numel()
    offsets = torch.empty(n_shards + 2, dtype=torch.int64, device=shard.device)
    perm = torch.empty(B, dtype=torch.int32, device=shard.device)
    pos = torch.empty(B, dtype=torch.int32, device=shard.device)
    check(lib.die_bucket_by_shard(ptr(shard), B, n_shards, ptr(offsets), ptr(perm), ptr(pos), stream_ptr()))
    return offsets, perm, pos


def pack_ragged(values: torch.Tensor, off: torch.Tensor, Lpad: int, src_index: Optional[torch.Tensor] = None,
                n_rows: Optional[int] = None):
    """values: 1-D tensor of payload elements, off: int64[n_req+1] -> (packed [R, Lpad], lengths int32[R])."""
    lib = require_cuda()
    R = (off.numel() - 1) if src_index is None else src_index.numel()
    if n_rows is not None:
        R = n_rows
    out = torch.empty((R, Lpad), dtype=values.dtype, device=values.device)
    lengths = torch.empty(R, dtype=torch.int32, device=values.device)
    check(lib.die_pack_ragged(ptr(values), ptr(off), ptr(src_index), R, values.element_size(), Lpad, ptr(out),
                              ptr(lengths), stream_ptr()))
    return out, lengths


def unpack_ragged(packed: torch.Tensor, off_out: torch.Tensor, out_values: torch.Tensor,
                  dst_index: Optional[torch.Tensor] = None):
    lib = require_cuda()
    R, Lpad = packed.shape
    check(lib.die_unpack_ragged(ptr(packed), ptr(off_out), ptr(dst_index), R, packed.element_size(), Lpad,
                                ptr(out_values), stream_ptr()))
    return out_values


def echo_fwd(x: torch.Tensor) -> torch.Tensor:
    lib = require_cuda()
    x = x.contiguous()
    y = torch.empty_like(x)
    check(lib.die_echo_fwd(ptr(x), ptr(y), x.numel() * x.element_size(), stream_ptr()))
    return y


def linear_fwd(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor], precision: int = _cabi.LINEAR_FP32):
    lib = require_cuda()
    assert x.dtype == torch.float32 and w.dtype == torch.float32 and x.is_contiguous() and w.is_contiguous()
    B, K = x.shape
    N = w.shape[0]
    y = torch.empty((B, N), dtype=torch.float32, device=x.device)
    check(lib.die_linear_fwd_f32(ptr(x), ptr(w), ptr(b), ptr(y), B, K, N, precision, stream_ptr()))
    return y


def mlp_fwd_bf16(x, w1, b1, w2, b2):
    lib = require_cuda()
    B, D = x.shape
    H = w1.shape[0]
    hidden = torch.empty((B, H), dtype=torch.bfloat16, device=x.device)
    y = torch.empty((B, D), dtype=torch.bfloat16, device=x.device)
    check(lib.die_mlp_fwd_bf16(ptr(x), ptr(w1), ptr(b1), ptr(w2), ptr(b2), ptr(hidden), ptr(y), B, D, H, stream_ptr()))
    return y


NAN = math.nan
