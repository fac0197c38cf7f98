"""Coordinator: the composition the reference's README promises and never wrote (README.md:56-60, SURVEY.md
call-stack (E)): Router -> per-shard KVStore -> Batcher pack -> Worker model -> unpack, for every shard of one
model version, on ONE GPU.  (One shard per GPU across a box is `die.exchange.PeerDispatcher`.)

Host code only: every stage is a kernel behind the C-ABI -- `die_route_md5` with the router's health masks
(src/router.py:140-221), `die_bucket_by_shard` (stable: request order inside a shard, which is the order the
reference's per-worker cache would see), `die_bucket_scatter_rows` / `die_bucket_gather_rows` for the hand-off, and one
`DispatchPipeline` per shard.  When the caches are at least a batch large (the serving case) nothing of a batch ever
comes back to the host: each shard's owner pass takes its request count from the device
(`die_pipeline_run_owner_device_n`) and runs on its own stream, so the shards' chains overlap on the GPU.  Tiny caches
(smaller than a batch) go through the general sequence with the counts read back, one shard after the other.
Requests the router returns None for (shard unhealthy and no healthy alternative) get a zero row, hit = 0 and
shard -1, as the reference's caller would see `None`.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch

from . import ops
from ._cabi import check, ptr, stream_ptr
from .pipeline import DispatchPipeline
from .router import Router


class Coordinator:
    def __init__(self, router: Router, model_name: str, version: str, kv_max_size: int, max_batch: int, dim_in: int,
                 dim_out: Optional[int] = None, model: str = "echo", weight=None, bias=None,
                 default_ttl: Optional[float] = None):
        mv = router.registry.get_model_version(model_name, version)
        if mv is None or not mv.shards:
            raise ValueError(f"Model {model_name}:{version} has no shards")
        self.router, self.model_name, self.version = router, model_name, version
        self.n_shards = len(mv.shards)
        self.dim_in, self.dim_out = dim_in, dim_out or dim_in
        # every shard owns its cache, exactly the reference's one-cache-per-worker model
        self.pipes: List[DispatchPipeline] = [
            DispatchPipeline(kv_max_size, max_batch, dim_in, self.dim_out, model, weight, bias, default_ttl)
            for _ in range(self.n_shards)]
        self.max_batch = max_batch
        self.on_device = max_batch <= kv_max_size  # the owner pass can take its batch size from the device
        if self.on_device:
            dev = torch.device("cuda", torch.cuda.current_device())
            S, C = self.n_shards, max_batch
            self._dig = torch.empty((S, C, 4), dtype=torch.int32, device=dev)
            self._pay = torch.empty((S, C, dim_in), dtype=torch.float32, device=dev)
            self._out = torch.empty((S, C, self.dim_out), dtype=torch.float32, device=dev)
            self._hit = torch.empty((S, C), dtype=torch.uint8, device=dev)
            self._cnt = torch.zeros(S, dtype=torch.int32, device=dev)
            self._streams = [torch.cuda.Stream() for _ in range(S)]

    def process(self, keys, payload: torch.Tensor, now: float,
                prefer_healthy: bool = True) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        """One flushed batch: keys (sequence of str or ops.KeyBatch), payload [B, dim_in] fp32 on the device.
        -> (out [B, dim_out] fp32, hit uint8 [B], shard int32 [B]) in request order."""
        kb = keys if isinstance(keys, ops.KeyBatch) else ops.KeyBatch.from_strings(keys)
        B = len(kb)
        dev = payload.device
        shard = self.router.route(self.model_name, self.version, kb, prefer_healthy)
        _, digests, _ = ops.route_md5(kb, 1)
        offsets, perm, pos = ops.bucket_by_shard(shard, self.n_shards)
        if self.on_device and B <= self.max_batch:
            return self._process_on_device(B, shard, digests, offsets, pos, payload.contiguous(), now)
        bounds = offsets.cpu().tolist()  # tiny caches: the general sequence sizes its launches on the host
        out = torch.zeros((B, self.dim_out), dtype=torch.float32, device=dev)
        hit = torch.zeros(B, dtype=torch.uint8, device=dev)
        for s in range(self.n_shards):
            lo, hi = bounds[s], bounds[s + 1]
            if hi == lo:
                continue
            idx = perm[lo:hi].long()
            o, h = self.pipes[s].run_owner_device(digests.index_select(0, idx).contiguous(),
                                                  payload.index_select(0, idx).contiguous(), now)
            out.index_copy_(0, idx, o)
            hit.index_copy_(0, idx, h)
        return out, hit, shard

    def _process_on_device(self, B, shard, digests, offsets, pos, payload, now):
        lib, S, C = self.pipes[0]._lib, self.n_shards, self.max_batch
        st = stream_ptr()
        check(lib.die_bucket_scatter_rows(ptr(digests), ptr(shard), ptr(pos), ptr(offsets), B, S, C, 16, ptr(self._dig),
                                          ptr(self._cnt), st))
        check(lib.die_bucket_scatter_rows(ptr(payload), ptr(shard), ptr(pos), ptr(offsets), B, S, C, self.dim_in * 4,
                                          ptr(self._pay), ptr(self._cnt), st))
        cur = torch.cuda.current_stream()
        for s in range(S):  # every shard's chain on its own stream: they overlap on the GPU
            side = self._streams[s]
            side.wait_stream(cur)
            check(lib.die_pipeline_run_owner_device_n(self.pipes[s]._h, ptr(self._dig[s]), C, ptr(self._cnt[s:]),
                                                      ptr(self._pay[s]), now, ptr(self._out[s]), ptr(self._hit[s]),
                                                      side.cuda_stream))
            cur.wait_stream(side)
        out = torch.empty((B, self.dim_out), dtype=torch.float32, device=payload.device)
        hit = torch.empty(B, dtype=torch.uint8, device=payload.device)
        check(lib.die_bucket_gather_rows(ptr(self._out), ptr(shard), ptr(pos), ptr(offsets), B, S, C, self.dim_out * 4,
                                         ptr(out), ptr(self._hit), ptr(hit), st))
        return out, hit, shard

    def stats(self) -> List[dict]:
        return [p.kv.stats() for p in self.pipes]
