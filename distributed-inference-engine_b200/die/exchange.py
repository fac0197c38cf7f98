"""Hash-sharded dispatch across ranks: shard list position s <-> rank s <-> GPU s.

Replaces the hop the reference makes over TCP + JSON, one request per connection
(src/worker.py:84-133), with two all-to-alls over NVLink per flushed batch:

  every rank ingests its own batch   -> route (K1, n_shards = world) -> stable bucket by owner
  dispatch  all-to-all               -> digests + payload rows travel to their owner rank
  owner     get -> pack -> model -> unpack -> set on its own KV shard (die_pipeline_run_owner_device);
            it sees the sources in rank order, request order kept inside each source
            (the order oracle/pipeline.py::process_exchange applies)
  return    all-to-all               -> results and hit flags travel back, scattered to request order

The collective calls go through torch.distributed (NCCL on GPUs); everything else is a libdie kernel.
The per-stage work is injected (`stages`) so that the exchange bookkeeping itself -- split sizes, ordering,
the two all-to-alls -- can be exercised with the gloo backend on CPU tensors by the test-suite.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist


class GpuStages:
    """The production stages: libdie kernels on the current CUDA device."""

    def __init__(self, world: int, kv_max_size: int, max_batch: int, dim: int, weight, bias, default_ttl,
                 model: str = "linear_tf32", dim_out: Optional[int] = None):
        from . import ops
        from .pipeline import DispatchPipeline
        self.ops = ops
        self.world = world
        self.dim, self.dim_out = dim, dim if dim_out is None else dim_out
        # an owner can receive more than one rank's batch worth of requests when the hash is unlucky
        self.pipe = DispatchPipeline(kv_max_size, 2 * max_batch, dim, self.dim_out, model, weight, bias, default_ttl)
        self.device = torch.device("cuda", torch.cuda.current_device())
        self._row_off = {}

    def row_offsets(self, n: int, width: int) -> torch.Tensor:
        key = (n, width)
        if key not in self._row_off:
            self._row_off[key] = torch.arange(n + 1, dtype=torch.int64, device=self.device) * width
        return self._row_off[key]

    def route(self, keys):
        shard, digests, _ = self.ops.route_md5(keys, self.world)
        return shard, digests

    def bucket(self, shard):
        return self.ops.bucket_by_shard(shard, self.world)

    def gather_rows(self, rows: torch.Tensor, perm: torch.Tensor) -> torch.Tensor:
        """rows[perm] with the pack kernel (fixed-width rows are a CSR with constant stride)."""
        flat = rows.reshape(-1)
        width = rows.shape[1]
        out, _ = self.ops.pack_ragged(flat, self.row_offsets(rows.shape[0], width), width, perm)
        return out

    def scatter_rows(self, rows: torch.Tensor, perm: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
        """out[perm[j]] = rows[j] with the unpack kernel."""
        self.ops.unpack_ragged(rows, self.row_offsets(out.shape[0], out.shape[1]), out.reshape(-1), perm)
        return out

    def owner(self, digests: torch.Tensor, payload: torch.Tensor, now: float):
        n = digests.shape[0]
        cap = self.pipe.max_batch
        if n <= cap:
            return self.pipe.run_owner_device(digests, payload, now)
        outs, hits = [], []
        for a in range(0, n, cap):  # rare: split, still in arrival order
            o, h = self.pipe.run_owner_device(digests[a:a + cap].contiguous(), payload[a:a + cap].contiguous(), now)
            outs.append(o)
            hits.append(h)
        return torch.cat(outs), torch.cat(hits)

    def empty(self, shape, dtype):
        return torch.empty(shape, dtype=dtype, device=self.device)


class ShardedDispatcher:
    def __init__(self, world: int, rank: int, kv_max_size: int = 0, max_batch: int = 0, dim: int = 0, weight=None,
                 bias=None, default_ttl=None, stages=None, group=None, model: str = "linear_tf32"):
        self.world, self.rank, self.group = world, rank, group
        self.stages = stages if stages is not None else GpuStages(world, kv_max_size, max_batch, dim, weight, bias,
                                                                  default_ttl, model)
        self.pipe = getattr(self.stages, "pipe", None)
        self.last_recv = 0

    def step(self, keys, payload: torch.Tensor, now: float, out: Optional[torch.Tensor] = None,
             hit: Optional[torch.Tensor] = None):
        """One flushed batch ingested on this rank.  Returns (out [B, dim_out], hit uint8 [B], shard int32 [B])."""
        st, W = self.stages, self.world
        B = payload.shape[0]
        shard, digests = st.route(keys)
        offsets, perm, _pos = st.bucket(shard)
        send_dig = st.gather_rows(digests, perm)
        send_pay = st.gather_rows(payload, perm)
        # split sizes: one small exchange + the step's only host read-back
        counts = (offsets[1:W + 1] - offsets[:W]).to(torch.int64)
        recv_counts = torch.empty_like(counts)
        dist.all_to_all_single(recv_counts, counts, group=self.group)
        in_split = counts.tolist()
        out_split = recv_counts.tolist()
        n_send, n_recv = sum(in_split), sum(out_split)
        self.last_recv = n_recv
        recv_dig = st.empty((n_recv, digests.shape[1]), digests.dtype)
        recv_pay = st.empty((n_recv, payload.shape[1]), payload.dtype)
        dist.all_to_all_single(recv_dig, send_dig[:n_send], out_split, in_split, group=self.group)
        dist.all_to_all_single(recv_pay, send_pay[:n_send], out_split, in_split, group=self.group)
        res, res_hit = st.owner(recv_dig, recv_pay, now) if n_recv else (
            st.empty((0, getattr(st, "dim_out", payload.shape[1])), payload.dtype), st.empty((0,), torch.uint8))
        back = st.empty((B, res.shape[1]), res.dtype)
        back_hit = st.empty((B, 1), torch.uint8)
        dist.all_to_all_single(back[:n_send], res, in_split, out_split, group=self.group)
        dist.all_to_all_single(back_hit[:n_send], res_hit.reshape(-1, 1), in_split, out_split, group=self.group)
        if n_send < B:  # requests the router returned None for: no result
            back[n_send:].zero_()
            back_hit[n_send:].zero_()
        out = st.empty((B, res.shape[1]), res.dtype) if out is None else out
        hit2 = st.empty((B, 1), torch.uint8)
        st.scatter_rows(back, perm, out)
        st.scatter_rows(back_hit, perm, hit2)
        if hit is not None:
            hit.copy_(hit2.reshape(-1))
            return out, hit, shard
        return out, hit2.reshape(-1), shard


def _set_health(lib, cabi, x, world, routable, healthy, failover):
    import ctypes
    if routable is None:
        cabi.check(lib.die_xchg_set_health(x, None, None, 0, 1 if failover else 0, cabi.stream_ptr()))
        return
    if len(routable) != world:
        raise ValueError("routable must have one entry per rank")
    r = (ctypes.c_uint8 * world)(*[1 if v else 0 for v in routable])
    healthy = list(healthy or [])
    h = (ctypes.c_int32 * max(len(healthy), 1))(*healthy)
    cabi.check(lib.die_xchg_set_health(x, ctypes.addressof(r), ctypes.addressof(h) if healthy else None, len(healthy),
                                       1 if failover else 0, cabi.stream_ptr()))


class PeerDispatcher:
    """The same dispatch over NVLink by direct peer stores (csrc/exchange.cu): no collective, no host sync.

    Every rank owns a receive window and a return window inside one symmetric allocation
    (torch.distributed._symmetric_memory); the dispatch kernel packs and writes each request row straight into
    its owner's window, a flag barrier inside the window separates the phases, the owner runs the fused
    pipeline on a batch whose size it only knows on the device, and the return / scatter kernels bring the
    results back to request order.  Owners see the sources in rank order (same contract as ShardedDispatcher).
    One `die_xchg_step` call enqueues the whole step (three in-order stage streams over four buffer sets inside
    the library, so consecutive steps overlap); `step_many` hands over a queue of batches, replayed in groups of
    16 steps from one CUDA graph.  `single_call=False` drives the stand-alone phase kernels from Python with
    torch's symmetric-memory barrier instead (used by the tests as a cross-check)."""

    def __init__(self, world: int, rank: int, kv_max_size: int, max_batch: int, dim: int, weight, bias, default_ttl,
                 model: str = "linear_tf32", group=None, single_call: bool = True, owner_max: int = 0):
        """owner_max: most requests an owner handles per step (default: two ingested batches' worth); a step that
        exceeds it raises in check().  Size it tightly when the model's cost grows with it (the bf16 MLP)."""
        import ctypes
        import torch.distributed._symmetric_memory as symm
        from . import _cabi, ops
        from .pipeline import DispatchPipeline
        self._cabi, self.ops, self._ct = _cabi, ops, ctypes
        self.lib = _cabi.require_cuda()
        self.world, self.rank, self.B, self.dim = world, rank, max_batch, dim
        self.single_call = single_call
        self.dev = torch.device("cuda", torch.cuda.current_device())
        self.cap = max_batch
        # hash sharding spreads every batch evenly over the owners: two batches' worth is far beyond any
        # fluctuation (sigma ~ sqrt(B)); a step that exceeds it raises in check() instead of being truncated silently
        self.bmax = (min(world, 2) if single_call else world) * self.cap
        if owner_max and single_call:
            self.bmax = min(self.bmax, int(owner_max))
        self.pipe = DispatchPipeline(kv_max_size, self.bmax, dim, dim, model, weight, bias, default_ttl)
        nbytes = int(self.lib.die_xchg_window_bytes(world, self.cap, dim))
        self.sym = symm.empty((nbytes + 3) // 4, dtype=torch.float32, device=self.dev)
        self.sym.zero_()
        torch.cuda.synchronize()
        self.hdl = symm.rendezvous(self.sym, group if group is not None else dist.group.WORLD)
        base = [int(p) for p in self.hdl.buffer_ptrs]
        arr = ctypes.c_void_p * world
        self.peer_windows = arr(*base)
        rb = world * self.cap * (dim + 4) * 4
        self.peer_rows = arr(*base)
        self.peer_back = arr(*[b + rb for b in base])
        self.peer_cnt = arr(*[b + 2 * rb for b in base])
        self.my_rows, self.my_back, self.my_cnt = base[rank], base[rank] + rb, base[rank] + 2 * rb
        h = ctypes.c_void_p()
        _cabi.check(self.lib.die_xchg_create(world, rank, self.cap, dim, self.bmax, self.pipe._h, self.peer_windows,
                                             ctypes.byref(h)))
        self._x = h.value
        self._prepared = set()
        e = lambda shape, dt: torch.empty(shape, dtype=dt, device=self.dev)  # noqa: E731
        self.recv_dig = e((self.bmax, 4), torch.int32)
        self.recv_pay = e((self.bmax, dim), torch.float32)
        self.recv_out = e((self.bmax, dim), torch.float32)
        self.recv_hit = e((self.bmax,), torch.uint8)
        self.n_dev = torch.zeros(1, dtype=torch.int32, device=self.dev)
        self.prefix = torch.zeros(world + 1, dtype=torch.int32, device=self.dev)
        torch.cuda.synchronize()
        self.hdl.barrier(channel=0)
        torch.cuda.synchronize()

    def __del__(self):
        if getattr(self, "_x", None):
            self.lib.die_xchg_destroy(self._x)
            self._x = None

    def step(self, keys, payload: torch.Tensor, now: float, out: Optional[torch.Tensor] = None,
             hit: Optional[torch.Tensor] = None):
        lib, check, ptr, st = self.lib, self._cabi.check, self._cabi.ptr, self._cabi.stream_ptr
        W, B, D = self.world, payload.shape[0], self.dim
        out = torch.empty((B, D), dtype=torch.float32, device=self.dev) if out is None else out
        hit = torch.empty(B, dtype=torch.uint8, device=self.dev) if hit is None else hit
        if self.single_call:
            shard = torch.empty(B, dtype=torch.int32, device=self.dev)
            check(lib.die_xchg_step(self._x, ptr(keys.key_bytes), ptr(keys.key_off), B, ptr(payload), now, ptr(out),
                                    ptr(hit), ptr(shard), st()))
            return out, hit, shard
        shard, digests, _ = self.ops.route_md5(keys, W)
        offsets, perm, _pos = self.ops.bucket_by_shard(shard, W)
        check(lib.die_xchg_dispatch(ptr(digests), ptr(payload), ptr(perm), ptr(offsets), B, D, W, self.rank, self.cap,
                                    self.peer_rows, self.peer_cnt, st()))
        self.hdl.barrier(channel=0)  # every rank's rows have landed in their owners' windows
        check(lib.die_xchg_compact(self.my_rows, self.my_cnt, W, self.cap, D, ptr(self.recv_dig), ptr(self.recv_pay),
                                   ptr(self.n_dev), ptr(self.prefix), st()))
        check(lib.die_pipeline_run_owner_device_n(self.pipe._h, ptr(self.recv_dig), self.bmax, ptr(self.n_dev),
                                                  ptr(self.recv_pay), now, ptr(self.recv_out), ptr(self.recv_hit), st()))
        check(lib.die_xchg_return(ptr(self.recv_out), ptr(self.recv_hit), ptr(self.prefix), W, self.rank, self.cap, D,
                                  self.peer_back, st()))
        self.hdl.barrier(channel=0)  # every owner's results are back in the sources' windows
        check(lib.die_xchg_scatter(self.my_back, ptr(perm), ptr(offsets), B, W, self.cap, D, ptr(out), ptr(hit), st()))
        return out, hit, shard

    def step_many(self, key_bytes: torch.Tensor, key_off: torch.Tensor, payload: torch.Tensor, now0: float,
                  now_step: float, out: torch.Tensor, hit: torch.Tensor) -> None:
        """A queue of flushed batches (key_bytes [n, stride], key_off [B+1] shared or [n, B+1], payload [n, B, dim]);
        out [ring, B, dim] / hit [ring, B] are reused round-robin.  Every rank must submit the same n."""
        n, B = payload.shape[0], payload.shape[1]
        shared = key_off.dim() == 1
        ptr = self._cabi.ptr
        if B not in self._prepared:  # capture + instantiate the replays before the first queue, not inside it
            self._cabi.check(self.lib.die_xchg_prepare(self._x, B))
            self._prepared.add(B)
        self._cabi.check(self.lib.die_xchg_step_many(self._x, n, ptr(key_bytes), key_bytes.stride(0), ptr(key_off),
                                                     0 if shared else key_off.stride(0), B, ptr(payload), now0, now_step,
                                                     ptr(out), ptr(hit), out.shape[0], self._cabi.stream_ptr()))

    def set_health(self, routable=None, healthy=None, failover: bool = True) -> None:
        """The router's view of the owners for the following steps: `Router.health_masks(shards, prefer_healthy)` gives
        both lists.  None: every rank routable.  Every rank must pass the same view."""
        _set_health(self.lib, self._cabi, self._x, self.world, routable, healthy, failover)

    def check(self) -> None:
        self._cabi.check(self.lib.die_xchg_check_sync(self._x, self._cabi.stream_ptr()))


class LoopbackPeerGroup:
    """All `world` ranks of the peer exchange inside ONE process on ONE GPU.

    Each rank gets its own owner pipeline (its own KV shard), its own `die_xchg_t` and its own window; the
    windows are ordinary device allocations of the same GPU, so "peer stores" land in local HBM.  The code that
    runs is exactly csrc/exchange.cu's sharded step -- route, dispatch, flag barrier, owner pass, return, scatter --
    which makes the N > 1 path checkable on a one-GPU box (tests/test_gpu_loopback.py, __graft_entry__.smoke()).

    The flag barriers of the ranks wait on one another, so every rank's call of a step must be enqueued before
    any of them is synchronised, the ranks' streams must not share a hardware queue (CUDA_DEVICE_MAX_CONNECTIONS=32
    before CUDA is initialised; tests/conftest.py and smoke() set it) and no kernel may be loaded for the first time
    while a barrier spins (the constructor runs a world-1 warm-up).  Table
    housekeeping synchronises one rank's streams: call `maintain()` between steps instead of letting a step
    trigger it while its peers' work is not enqueued yet."""

    def __init__(self, world: int, kv_max_size: int, max_batch: int, dim: int, weight, bias, default_ttl,
                 model: str = "linear_tf32", _warm: bool = True):
        import ctypes
        from . import _cabi, ops
        from .pipeline import DispatchPipeline
        self._cabi, self.ops = _cabi, ops
        self.lib = _cabi.require_cuda()
        self.world, self.cap, self.dim = world, max_batch, dim
        self.dev = torch.device("cuda", torch.cuda.current_device())
        self.bmax = min(world, 2) * max_batch
        self.pipes = [DispatchPipeline(kv_max_size, self.bmax, dim, dim, model, weight, bias, default_ttl)
                      for _ in range(world)]
        nbytes = int(self.lib.die_xchg_window_bytes(world, self.cap, dim))
        self.windows = [torch.zeros((nbytes + 3) // 4, dtype=torch.float32, device=self.dev) for _ in range(world)]
        self._peers = (ctypes.c_void_p * world)(*[w.data_ptr() for w in self.windows])
        self._x = []
        # All ranks' stage streams live on ONE device here and a barrier kernel spins on its stream's hardware queue until
        # its peers arrive: the group's streams (2 * ways + 1 per rank) must fit the device's 32 queues with room for the
        # callers' streams, or a peer's kernel can sit behind the very barrier that waits for it.
        import os
        ways_before = os.environ.get("DIE_XCHG_WAYS")
        if ways_before is None:
            os.environ["DIE_XCHG_WAYS"] = str(max(1, min(4, (24 // world - 1) // 2)))
        try:
            for r in range(world):
                h = ctypes.c_void_p()
                _cabi.check(self.lib.die_xchg_create(world, r, self.cap, dim, self.bmax, self.pipes[r]._h, self._peers,
                                                     ctypes.byref(h)))
                self._x.append(h.value)
        finally:
            if ways_before is None:
                del os.environ["DIE_XCHG_WAYS"]
        self.streams = [torch.cuda.Stream() for _ in range(world)]
        torch.cuda.synchronize()
        if _warm:
            # CUDA loads a kernel at its first launch, and that load synchronises the context: with a peer's barrier
            # already spinning on this device it would wait for the barrier's time-out.  A world-1 group (whose
            # barriers wait for nobody) launches every kernel of the step once, stream-launched and replayed.
            warm = LoopbackPeerGroup(1, kv_max_size, max_batch, dim, weight, bias, default_ttl, model, _warm=False)
            kb = self.ops.KeyBatch.from_strings(["warm-%d" % i for i in range(max_batch)])
            x = torch.zeros((max_batch, dim), dtype=torch.float32, device=self.dev)
            warm.step([kb], [x], 0.0)
            n = max(2, warm.pipes[0]._graph_steps())
            warm.step_many([kb.key_bytes.repeat(n, 1)], [kb.key_off.repeat(n, 1)], [x.repeat(n, 1, 1)], 1.0, 1.0)
            del warm
            torch.cuda.synchronize()

    def __del__(self):
        for h in getattr(self, "_x", []):
            self.lib.die_xchg_destroy(h)
        self._x = []

    def _sync_and_check(self):
        for s in self.streams:
            s.synchronize()
        torch.cuda.synchronize()
        for r, h in enumerate(self._x):
            self._cabi.check(self.lib.die_xchg_check_sync(h, self.streams[r].cuda_stream))

    def maintain(self):
        for p in self.pipes:
            p.kv.maintain(force=False)

    def set_health(self, routable=None, healthy=None, failover: bool = True) -> None:
        for r, h in enumerate(self._x):
            with torch.cuda.stream(self.streams[r]):
                _set_health(self.lib, self._cabi, h, self.world, routable, healthy, failover)

    def step(self, keys, payloads, now: float):
        """keys[r] / payloads[r]: the batch rank r ingests.  -> [(out, hit, shard)] per rank."""
        check, ptr = self._cabi.check, self._cabi.ptr
        # every allocation first: a cudaMalloc synchronises the device and would wait for a peer's spinning barrier
        res = [(torch.empty((p.shape[0], self.dim), dtype=torch.float32, device=self.dev),
                torch.empty(p.shape[0], dtype=torch.uint8, device=self.dev),
                torch.empty(p.shape[0], dtype=torch.int32, device=self.dev)) for p in payloads]
        torch.cuda.synchronize()
        for r in range(self.world):
            out, hit, shard = res[r]
            check(self.lib.die_xchg_step(self._x[r], ptr(keys[r].key_bytes), ptr(keys[r].key_off), payloads[r].shape[0],
                                         ptr(payloads[r]), now, ptr(out), ptr(hit), ptr(shard), self.streams[r].cuda_stream))
        self._sync_and_check()
        return res

    def step_many(self, key_bytes, key_off, payloads, now0: float, now_step: float):
        """key_bytes[r] [n, stride], key_off[r] [n, B+1], payloads[r] [n, B, dim] -> [(out [n,B,dim], hit [n,B])]."""
        check, ptr = self._cabi.check, self._cabi.ptr
        bufs = []
        for r in range(self.world):   # allocations and graph instantiation before any rank's barriers are enqueued
            n, B = payloads[r].shape[0], payloads[r].shape[1]
            bufs.append((torch.empty((max(n, 2), B, self.dim), dtype=torch.float32, device=self.dev),
                         torch.empty((max(n, 2), B), dtype=torch.uint8, device=self.dev)))
            check(self.lib.die_xchg_prepare(self._x[r], B))
        torch.cuda.synchronize()
        res = []
        for r in range(self.world):
            n, B = payloads[r].shape[0], payloads[r].shape[1]
            out, hit = bufs[r]
            check(self.lib.die_xchg_step_many(self._x[r], n, ptr(key_bytes[r]), key_bytes[r].stride(0), ptr(key_off[r]),
                                              key_off[r].stride(0), B, ptr(payloads[r]), now0, now_step, ptr(out), ptr(hit),
                                              out.shape[0], self.streams[r].cuda_stream))
            res.append((out[:n], hit[:n]))
        self._sync_and_check()
        return res
