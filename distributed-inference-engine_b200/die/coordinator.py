"""Coordinator: the composition the reference's README promises and never wrote (README.md:56-60, SURVEY.md
call-stack (E)): Router -> per-shard KVStore -> Batcher pack -> Worker model -> unpack, for every shard of one
model version, on ONE GPU.  (One shard per GPU across a box is `die.exchange.PeerDispatcher`.)

Host code only: every stage is a kernel behind the C-ABI -- `die_route_md5` with the router's health masks
(src/router.py:140-221), `die_bucket_by_shard` (stable: request order inside a shard, which is the order the
reference's per-worker cache would see), and one `DispatchPipeline` per shard (`die_pipeline_run_owner_device`).
torch indexing is used for the gathers between them (plumbing).  Requests the router returns None for
(shard unhealthy and no healthy alternative) get a zero row, hit = 0 and shard -1, as the reference's caller
would see `None`.
"""
from __future__ import annotations

from typing import List, Optional, Sequence, Tuple

import torch

from . import ops
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

    def process(self, keys, payload: torch.Tensor, now: float,
                prefer_healthy: bool = True) -> Tuple[torch.Tensor, torch.Tensor, torch.Tensor]:
        """One flushed batch: keys (sequence of str or ops.KeyBatch), payload [B, dim_in] fp32 on the device.
        -> (out [B, dim_out] fp32, hit uint8 [B], shard int32 [B]) in request order."""
        kb = keys if isinstance(keys, ops.KeyBatch) else ops.KeyBatch.from_strings(keys)
        B = len(kb)
        dev = payload.device
        shard = self.router.route(self.model_name, self.version, kb, prefer_healthy)
        _, digests, _ = ops.route_md5(kb, 1)
        offsets, perm, _ = ops.bucket_by_shard(shard, self.n_shards)
        bounds = offsets.cpu().tolist()  # the per-shard batch sizes are needed on the host to size the launches
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

    def stats(self) -> List[dict]:
        return [p.kv.stats() for p in self.pipes]
