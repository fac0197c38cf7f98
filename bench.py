#!/usr/bin/env python
"""Headline benchmark: requests/sec of the batched route -> probe -> pack -> infer -> unpack path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json configs[1], "C2"): flushed batches of 4096 requests, key "user-%010d" with ids
uniform in [0, 1e7), payload 128 fp32, mock linear model 128->128, one owner KV cache of 2^20 entries
per GPU (LRU, default_ttl 300 s), so about 10 % of the requests hit the cache and every batch evicts.
One step = one flushed batch through the whole path.  Every step has its own freshly drawn inputs
(resident in HBM before the timed region; K steps x 2.2 MB, larger than L2 for K >= 60).

Prints ONE JSON line (see the contract in the task description): `value` is device-resident
throughput, `e2e` goes through the C-ABI with host buffers, `roofline` describes the dominant kernel
timed with CUDA events on its own stream, `cpu_baseline` is the CPU oracle port timed on this host.
"""
import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "distributed-inference-engine_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

B, D = 4096, 128
KV_MAX, ID_SPACE, TTL = 1 << 20, 10 ** 7, 300.0
KEY_LEN = 15
WORLD = int(os.environ.get("WORLD_SIZE", "1"))
METRIC = "requests/sec, batched route->probe->pack->infer->unpack"
MODEL = os.environ.get("DIE_BENCH_MODEL", "linear_tf32")  # tcgen05 kind::tf32 GEMM, fp32 accumulate


# ----------------------------------------------------------------------------------------------------------------
# synthetic inputs
# ----------------------------------------------------------------------------------------------------------------
def key_bytes_of(ids: np.ndarray) -> np.ndarray:
    """b"user-%010d" for every id -> uint8 [n, 15] without a Python loop."""
    digits = (ids[:, None] // (10 ** np.arange(9, -1, -1, dtype=np.int64))[None, :]) % 10 + 48
    out = np.empty((ids.size, KEY_LEN), dtype=np.uint8)
    out[:, :5] = np.frombuffer(b"user-", dtype=np.uint8)
    out[:, 5:] = digits.astype(np.uint8)
    return out


def make_batches(rng, n_batches, id_space=None, payload=True):
    id_space = id_space or ID_SPACE * WORLD  # weak scaling: the key population grows with the number of shards
    ids = rng.integers(0, id_space, (n_batches, B), dtype=np.int64)
    keys = key_bytes_of(ids.reshape(-1)).reshape(n_batches, B * KEY_LEN)
    off = np.arange(B + 1, dtype=np.int64) * KEY_LEN
    x = rng.standard_normal((n_batches, B, D), dtype=np.float32) if payload else None
    return ids, keys, off, x


def model_weights():
    rng = np.random.default_rng(12345)
    w = (rng.standard_normal((D, D)) / np.sqrt(D)).astype(np.float32)
    b = rng.standard_normal(D).astype(np.float32)
    return w, b


# ----------------------------------------------------------------------------------------------------------------
# CPU baseline (oracle port of the reference's pure-Python path) -- the one place bench.py may run oracle/
# ----------------------------------------------------------------------------------------------------------------
def cpu_worker(seed, n_timed, scale=8):
    """Prefill a scaled-down cache (1/scale of the entries and of the id space: same ~10 % hit rate),
    then time n_timed flushed batches through the composed reference port.  Returns (requests, seconds)."""
    from oracle.pipeline import BaselinePipeline
    w, b = model_weights()
    w64 = w.astype(np.float64)

    def model(rows):
        if not rows:
            return []
        y = (np.stack(rows) @ w.T + b).astype(np.float32)
        return [y[i] for i in range(len(rows))]

    kv_max, space = KV_MAX // scale, ID_SPACE // scale
    pipe = BaselinePipeline(1, kv_max, TTL, model, max_batch_size=B)
    rng = np.random.default_rng(1000 + seed)
    n_fill = int(kv_max * 1.15 / B) + 1
    now = 1000.0
    for _ in range(n_fill):
        ids = rng.integers(0, space, B)
        keys = ["user-%010d" % i for i in ids]
        x = rng.standard_normal((B, D), dtype=np.float32)
        now += 0.01
        pipe.process(keys, list(x), now)
    batches = []
    for _ in range(n_timed):
        ids = rng.integers(0, space, B)
        batches.append((["user-%010d" % i for i in ids], list(rng.standard_normal((B, D), dtype=np.float32))))
    t0 = time.perf_counter()
    for keys, rows in batches:
        now += 0.01
        pipe.process(keys, rows, now)
    dt = time.perf_counter() - t0
    del w64
    return n_timed * B, dt


def cpu_stage_times(n_rep=3):
    """Per-stage time of the reference's path on one host core (oracle port), one 4096-request batch each:
    the CPU column beside the per-kernel GPU numbers."""
    import hashlib
    from oracle.ref_port import PortKVCache, PortRegistry
    rng = np.random.default_rng(4242)
    w, b = model_weights()
    reg = PortRegistry()
    reg.register_model("m", "v1")
    for sid in range(8):
        reg.add_shard("m", "v1", sid, f"w{sid}")
    clock = [1000.0]
    cache = PortKVCache(KV_MAX // 8, default_ttl=TTL, clock=lambda: clock[0])
    keys = ["user-%010d" % i for i in rng.integers(0, ID_SPACE // 8, B)]
    for k in ["user-%010d" % i for i in rng.integers(0, ID_SPACE // 8, 20 * B)]:
        cache.set(k, 0)
    rows = list(rng.standard_normal((B, D), dtype=np.float32))
    best = {}

    def timed(name, fn):
        t = min(_time_once(fn) for _ in range(n_rep))
        best[name] = {"us_per_request": round(1e6 * t / B, 3), "requests_per_s": round(B / t, 1)}

    timed("route (md5 mod N, model_registry.py:159)", lambda: [reg.shard_for_key("m", "v1", k) for k in keys])
    timed("probe (KVCache.get, kvstore.py:136)", lambda: [cache.get(k) for k in keys])
    timed("fill (KVCache.set + LRU eviction, kvstore.py:104)", lambda: [cache.set(k + "x", 1) for k in keys])
    timed("pack + unpack (list gather / zip, batcher.py:211,225)", lambda: [r for r, _ in zip([q for q in rows], rows)])
    timed("infer (numpy fp32 linear on the packed batch)", lambda: (np.stack(rows) @ w.T + b))
    del hashlib
    return best


def _time_once(fn):
    t0 = time.perf_counter()
    fn()
    return time.perf_counter() - t0


def cpu_baseline_single(n_timed=24):
    n, dt = cpu_worker(0, n_timed)
    per_batch = dt / n_timed
    try:
        stages = cpu_stage_times()
    except Exception as exc:  # noqa: BLE001
        stages = {"error": repr(exc)[:200]}
    return {"value": n / dt, "unit": "requests/s", "cores": 1, "kind": "port",
            "as_shipped": {"value": B / (per_batch + 0.1), "unit": "requests/s",
                           "note": "the same path with the reference's own sleep left in: mock_batch_inference waits "
                                   "100 ms per batch (mock_inference.py:12-73), added analytically to the measured batch time"},
            "stages": stages,
            "sample": f"{n_timed} batches x {B} requests after prefill; cache 2^17 entries and id space 1.25e6 "
                      f"(both 1/8 of the GPU run, same ~10% hit rate); oracle/pipeline.py BaselinePipeline "
                      f"(hashlib.md5 + dict/OrderedDict + asyncio batcher + numpy fp32 linear), sleeps removed"}


def run_reference(args):
    """`--impl reference`: the reference's pure-Python path (oracle port; the Python reference itself cannot
    travel to the GPU box) on all usable host cores, as independent one-shard replicas."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    cores = max(1, min(os.cpu_count() or 1, 32))
    n_timed = max(4, min(args.steps, 24))
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(cores) as pool:
        res = pool.starmap(cpu_worker, [(s, n_timed) for s in range(cores)])
    total = sum(r[0] for r in res)
    slowest = max(r[1] for r in res)
    value = total / slowest
    line = {"metric": METRIC, "value": value, "unit": "requests/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * slowest / n_timed, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": workload_config(args.gpus),
            "cpu_baseline": {"value": value, "unit": "requests/s", "cores": cores, "kind": "port",
                             "sample": f"{cores} replica processes x {n_timed} batches x {B} requests after prefill, "
                                       f"cache and id space scaled 1/8; wall {time.perf_counter() - t0:.1f}s"},
            "e2e": {"value": value, "unit": "requests/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def workload_config(n_gpus):
    return {"workload": "C2: 4096-request flushed batches per GPU, key 'user-%010d' (ids uniform in [0,1e7 x n_gpus)), "
                        "128-dim fp32 payload, mock linear 128->128, LRU+TTL cache of 2^20 entries per GPU "
                        "(~10% hits, evicting)",
            "batch": B, "dim": D, "kv_max_size": KV_MAX, "id_space": ID_SPACE * n_gpus, "default_ttl": TTL,
            "parallelism": f"hash-shard x{n_gpus}", "l2_policy": "every step reads fresh inputs; K steps x 2.2 MB"}


# ----------------------------------------------------------------------------------------------------------------
# clocks
# ----------------------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """SM clock and throttle reasons while the timed region executes (pynvml = what nvidia-smi reads).

    An NVML query is not free here: it takes 2-3 ms and, on these boxes, CUDA-graph launches issued meanwhile
    wait for it (a 2 ms polling period slowed the step 6x).  So the sampler never queries while the host is
    still submitting work: the bench calls sample_now() right after the last step of the timed region is
    enqueued, i.e. while the GPU is still executing that region, and again every `period` (the recipe's 200 ms)
    only if the region lasts that long."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.samples, self.reasons, self.stop_flag, self.max_mhz = [], set(), False, None
        self.period = float(os.environ.get("DIE_BENCH_SAMPLER_MS", "200")) / 1e3
        self.query_s = 0.0
        self.wake = threading.Event()
        self.requested = self.served = 0
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:  # noqa: BLE001
            self.nv = None

    def _query(self):
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4)}
        try:
            t_q = time.perf_counter()
            self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
            try:
                mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
            except Exception:  # noqa: BLE001
                mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
            for k, bit in names.items():
                if mask & bit:
                    self.reasons.add(k)
            self.query_s += time.perf_counter() - t_q
        except Exception:  # noqa: BLE001
            pass

    def run(self):
        if self.nv is None or self.period <= 0:
            return
        while not self.stop_flag:
            self.wake.wait(timeout=self.period)
            if self.stop_flag:
                break
            self.wake.clear()
            want = self.requested
            self._query()           # periodic (timeout) or on demand (sample_now)
            self.served = max(self.served, want)

    def sample_now(self):
        self.requested += 1
        self.wake.set()

    def result(self):
        t_end = time.perf_counter() + 1.0
        while self.nv is not None and self.period > 0 and self.served < self.requested and time.perf_counter() < t_end:
            time.sleep(0.0005)   # the requested query is still in flight
        self.stop_flag = True
        self.wake.set()
        self.join(timeout=1.0)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s), "period_ms": 1e3 * self.period,
                "nvml_ms_per_sample": round(1e3 * self.query_s / len(s), 3) if s else None,
                "when": "after the last step of the timed region was enqueued, while the GPU was still executing it"}


def ncu_traffic(kernel_name):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel_name`, from the committed
    `ncu --set full` capture of this bench command (profiles/ncu_full_r01_pipeline_kernels.csv)."""
    import csv
    path = os.path.join(ROOT, "profiles", "ncu_full_r01_pipeline_kernels.csv")
    try:
        rows = list(csv.reader(open(path)))
        hdr, units = rows[0], rows[1]
        i_name, i_r, i_w = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        vals = [float(r[i_r]) * scale[units[i_r]] + float(r[i_w]) * scale[units[i_w]]
                for r in rows[2:] if len(r) == len(hdr) and kernel_name in r[i_name]]
        if vals:
            return round(sum(vals) / len(vals)), "profiles/ncu_full_r01_pipeline_kernels.csv (mean of %d launches)" % len(vals)
    except Exception:  # noqa: BLE001
        pass
    return None, None


def e2e_sharded(runner, rng, K, now, off_d, world):
    """End to end at N > 1: every rank's batches start in pinned host memory and its results end there.  Groups of
    16 batches: one H2D copy per array, one die_xchg_step_many call (one graph replay), one D2H copy per array;
    three device staging sets on three streams keep the copy engines on the neighbouring groups.  Wall clock
    between two barriers, max over ranks."""
    import torch
    import torch.distributed as dist
    G = 16
    n_groups = max(2, K // G)
    n = n_groups * G
    _, keys, _, x = make_batches(rng, n)
    hk = torch.from_numpy(keys).pin_memory()
    hx = torch.from_numpy(x).pin_memory()
    hout = torch.empty((n, B, D), dtype=torch.float32).pin_memory()
    hhit = torch.empty((n, B), dtype=torch.uint8).pin_memory()
    NS = 3
    dk = [torch.zeros((G + 1, B * KEY_LEN), dtype=torch.uint8, device="cuda") for _ in range(NS)]  # +1 row of slack
    dx = [torch.empty((G, B, D), dtype=torch.float32, device="cuda") for _ in range(NS)]
    dout = [torch.empty((G, B, D), dtype=torch.float32, device="cuda") for _ in range(NS)]
    dhit = [torch.empty((G, B), dtype=torch.uint8, device="cuda") for _ in range(NS)]
    s_in, s_out, cur = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.current_stream()
    ev_in = [torch.cuda.Event() for _ in range(NS)]
    ev_done = [torch.cuda.Event() for _ in range(NS)]
    ev_out = [torch.cuda.Event() for _ in range(NS)]

    def run(first_group, count, now0):
        for g in range(first_group, first_group + count):
            b = g % NS
            lo, hi = (g % n_groups) * G, (g % n_groups) * G + G
            with torch.cuda.stream(s_in):
                if g - first_group >= NS:
                    s_in.wait_event(ev_done[b])
                dk[b][:G].copy_(hk[lo:hi], non_blocking=True)
                dx[b].copy_(hx[lo:hi], non_blocking=True)
                ev_in[b].record(s_in)
            cur.wait_event(ev_in[b])
            if g - first_group >= NS:
                cur.wait_event(ev_out[b])
            runner.step_many(dk[b][:G], off_d, dx[b], now0 + (g - first_group) * G * DT_E2E, DT_E2E, dout[b], dhit[b])
            ev_done[b].record(cur)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_done[b])
                hout[lo:hi].copy_(dout[b], non_blocking=True)
                hhit[lo:hi].copy_(dhit[b], non_blocking=True)
                ev_out[b].record(s_out)
        s_out.synchronize()
        cur.synchronize()

    run(0, 2, now)                         # warm the staging path and the graphs
    now += 2 * G * DT_E2E
    torch.cuda.synchronize()
    dist.barrier()
    t0 = time.perf_counter()
    run(0, n_groups, now)
    dt = time.perf_counter() - t0
    t = torch.tensor([dt], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dt = float(t[0])
    runner.check()
    return {"value": world * n * B / dt, "unit": "requests/s",
            "h2d_bytes_per_step": B * KEY_LEN + B * D * 4, "d2h_bytes_per_step": B * D * 4 + B,
            "ms_per_step": 1e3 * dt / n, "steps": n,
            "api": "PeerDispatcher.step_many (die_xchg_step_many) on groups of 16 batches staged from / to pinned host "
                   "buffers on every rank; copy-in, sharded steps and copy-out of consecutive groups overlap"}, now + n * DT_E2E


DT_E2E = 0.01


# ----------------------------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=2000)
    ap.add_argument("--warmup", type=int, default=50)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-kernels", action="store_true",
                    help="skip the per-kernel rooflines at C3 / C4 sizes (tools/bench_kernels.py, about 30 s)")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from die import _cabi, ops
    from die.pipeline import DispatchPipeline

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    lib = _cabi.require_cuda()
    K, W = args.steps, max(args.warmup, 3)
    rng = np.random.default_rng(7 + rank)
    w, b = model_weights()

    if world > 1:
        from die.exchange import PeerDispatcher, ShardedDispatcher
        cls = ShardedDispatcher if os.environ.get("DIE_BENCH_EXCHANGE", "peer") == "nccl" else PeerDispatcher
        runner = cls(world, rank, KV_MAX, B, D, w, b, TTL, model=MODEL)
    else:
        runner = None
    pipe = DispatchPipeline(KV_MAX, B * (2 if world > 1 else 1), D, D, MODEL, w, b, TTL, max_key_bytes=2 * B * KEY_LEN) \
        if runner is None else runner.pipe

    off_d = torch.from_numpy(np.arange(B + 1, dtype=np.int64) * KEY_LEN).cuda()
    RING = 64
    out_ring = torch.empty((RING, B, D), dtype=torch.float32, device="cuda")
    hit_ring = torch.empty((RING, B), dtype=torch.uint8, device="cuda")
    DT = 0.01  # seconds of injected clock per flushed batch

    def to_device(keys, x):
        return torch.from_numpy(keys).cuda(), torch.from_numpy(x).cuda()

    def run_steps(kd, xd, now0):
        """kd [n, B*15] uint8, xd [n, B, D]: n flushed batches, applied in order."""
        n = kd.shape[0]
        if runner is None:  # one C-ABI call submits the whole queue (each batch is still its own pass)
            pipe.run_device_many(kd, off_d, xd, now0, DT, out_ring, hit_ring)
        elif hasattr(runner, "step_many"):
            runner.step_many(kd, off_d, xd, now0, DT, out_ring, hit_ring)
        else:
            for i in range(n):
                runner.step(ops.KeyBatch(kd[i], off_d), xd[i], now0 + i * DT, out_ring[i % RING], hit_ring[i % RING])
        return now0 + n * DT

    # ---- prefill the cache to capacity (untimed) -------------------------------------------------------------
    now = 1000.0
    n_fill = int(KV_MAX * 1.15 / B) + 1
    for c in range(0, n_fill, 64):
        _, keys, _, x = make_batches(rng, min(64, n_fill - c))
        now = run_steps(*to_device(keys, x), now)
    torch.cuda.synchronize()

    # ---- timed: inputs resident in HBM ------------------------------------------------------------------------
    _, keys, _, x = make_batches(rng, W + K)
    kd, xd = to_device(keys, x)
    now = run_steps(kd[:W], xd[:W], now)
    torch.cuda.synchronize()
    stats0 = pipe.kv.stats()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = lib.die_launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    t_host = time.perf_counter()
    now = run_steps(kd[W:], xd[W:], now)
    t_host = time.perf_counter() - t_host  # host time to enqueue the queue (no wait on the GPU)
    ev1.record()
    sampler.sample_now()  # everything is enqueued: the query can no longer hold up a launch
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = ev0.elapsed_time(ev1)
    launches = lib.die_launch_count() - launches0
    clocks = sampler.result()
    stats1 = pipe.kv.stats()
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t[0])
    value = world * K * B / (ms / 1e3)
    hit_rate = (stats1["hits"] - stats0["hits"]) / max(1, (stats1["hits"] - stats0["hits"]) + (stats1["misses"] - stats0["misses"]))

    # ---- e2e: host buffers through the C-ABI (H2D + kernels + D2H inside the timed region) ---------------------
    e2e = None
    if world > 1 and hasattr(runner, "step_many"):
        try:
            e2e, now = e2e_sharded(runner, rng, K, now, off_d, world)
        except Exception as exc:  # noqa: BLE001  (the device-resident number above stands on its own)
            e2e = {"value": None, "unit": "requests/s", "error": repr(exc)[:300]}
    if world == 1:
        _, keys, off, x = make_batches(rng, K + 3)
        hk = torch.from_numpy(np.concatenate([keys, np.zeros((K + 3, 16), np.uint8)], axis=1)).pin_memory()
        ho = torch.from_numpy(off).pin_memory()
        hx = torch.from_numpy(x).pin_memory()
        hout = torch.empty((K, B, D), dtype=torch.float32).pin_memory()
        hhit = torch.empty((K, B), dtype=torch.uint8).pin_memory()
        pipe.run_host_many(hk[K:], ho, hx[K:], now, DT, hout[:3], hhit[:3])  # warm the staging path
        now += 3 * DT
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        pipe.run_host_many(hk[:K], ho, hx[:K], now, DT, hout, hhit)   # returns after the last result is on the host
        dt = time.perf_counter() - t0
        now += K * DT
        # the unpipelined form of the same call, one synchronous batch at a time
        t0 = time.perf_counter()
        n_sync = min(K, 50)
        for i in range(n_sync):
            pipe.run_host(hk[i], ho, hx[i], now + i * DT, hout[i], hhit[i])
        dt_sync = time.perf_counter() - t0
        now += n_sync * DT
        e2e = {"value": K * B / dt, "unit": "requests/s",
               "h2d_bytes_per_step": B * KEY_LEN + (B + 1) * 8 + B * D * 4, "d2h_bytes_per_step": B * D * 4 + B,
               "ms_per_step": 1e3 * dt / K,
               "api": "die_pipeline_run_host_many (C-ABI, pinned host buffers; copy-in / kernels / copy-out of "
                      "consecutive batches overlap)",
               "per_call_sync": {"value": n_sync * B / dt_sync, "ms_per_step": 1e3 * dt_sync / n_sync,
                                 "api": "die_pipeline_run_host, one blocking call per batch"}}

    # ---- per-stage device time (CUDA events on the launching stream) and roofline -------------------------------
    stages, roofline = {}, None
    if world == 1 or os.environ.get("DIE_BENCH_PROFILE_ALL"):
        n_prof = min(K, 50)
        _, keys, _, x = make_batches(rng, n_prof)
        kd2, xd2 = to_device(keys, x)
        s0 = pipe.kv.stats()
        _cabi.check(lib.die_profile_begin())
        now = run_steps(kd2, xd2, now)
        import ctypes
        ms_arr, cnt_arr, n_st = (ctypes.c_double * 32)(), (ctypes.c_uint64 * 32)(), ctypes.c_int32()
        _cabi.check(lib.die_profile_end(32, ms_arr, cnt_arr, ctypes.byref(n_st)))
        s1 = pipe.kv.stats()
        hits = (s1["hits"] - s0["hits"]) / n_prof
        miss = (s1["misses"] - s0["misses"]) / n_prof
        evic = (s1["evictions"] - s0["evictions"]) / n_prof
        row = D * 4
        alg = {  # algorithmic bytes per launch (SURVEY.md section 8(d) per-request figures x requests per launch)
            "kv_get": B * (KEY_LEN + 8 + 16) + hits * 121 + miss * 89 + B * 5 + evic * 12,   # route + probe + rank + window
            "kv_set_claim": miss * 160 + B * 2 * row + B * 13,                                # insert + pack / hit copy
            "model": miss * (2 * row + 16),
            "unpack_store": miss * 3 * row + B * 9 + evic * 128,                              # unpack + cached row + evict
        }
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:  # noqa: BLE001
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        peak_src = "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
        total_ms = sum(ms_arr[i] for i in range(n_st.value))
        kernel_of = {"kv_get": "pipe_get_kernel", "kv_set_claim": "pipe_claim_pack_kernel", "model": "linear_tf32_kernel",
                     "unpack_store": "pipe_unpack_evict_kernel"}
        for i in range(n_st.value):
            if cnt_arr[i] == 0:
                continue
            name = lib.die_profile_stage_name(i).decode()
            per = ms_arr[i] / cnt_arr[i]
            gbs = alg.get(name, 0) / (per * 1e-3) / 1e9 if per > 0 else 0.0
            stages[name] = {"kernel": kernel_of.get(name, name), "us_per_launch": round(per * 1e3, 3),
                            "launches_per_step": cnt_arr[i] / n_prof, "share": round(ms_arr[i] / total_ms, 4),
                            "alg_bytes_per_launch": round(alg.get(name, 0)), "gbs": round(gbs, 1),
                            "frac_of_hbm_peak": round(gbs / peak, 4)}
        dom = max(stages, key=lambda n: stages[n]["share"])
        traffic, traffic_src = ncu_traffic(stages[dom]["kernel"])
        roofline = {"kernel": stages[dom]["kernel"], "bound": "hbm", "achieved": stages[dom]["gbs"], "peak": peak,
                    "unit": "GB/s", "frac": stages[dom]["frac_of_hbm_peak"], "traffic": traffic, "traffic_source": traffic_src,
                    "peak_source": peak_src,
                    "us_per_launch": stages[dom]["us_per_launch"], "alg_bytes_per_launch": stages[dom]["alg_bytes_per_launch"],
                    "note": "a 4096-request batch carries ~5 MB of algorithmic traffic (<1 us at peak): every kernel of the "
                            "step sits on its launch + dependent-DRAM-latency floor, see DESIGN.md section 5"}

    # ---- host-side API edge of the drop-in classes: str -> bytes -> device, list of rows -> padded device batch ----
    api_edge = None
    if world == 1:
        try:
            from die.batcher import pack_requests
            ids1, _, _, x1 = make_batches(rng, 1)
            strs = ["user-%010d" % i for i in ids1[0]]
            rows = [x1[0][i] for i in range(B)]
            ops.KeyBatch.from_strings(strs)
            pack_requests(rows, Lpad=D)
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(5):
                ops.KeyBatch.from_strings(strs)
            torch.cuda.synchronize()
            t_keys = (time.perf_counter() - t0) / 5
            t0 = time.perf_counter()
            for _ in range(5):
                pack_requests(rows, Lpad=D)
            torch.cuda.synchronize()
            t_pack = (time.perf_counter() - t0) / 5
            api_edge = {"keys_str_to_device_us_per_batch": round(1e6 * t_keys, 1),
                        "payload_rows_to_packed_device_us_per_batch": round(1e6 * t_pack, 1),
                        "note": "Python-object edge of Router.route(keys) / Batcher (encode + H2D + pack kernel) for one "
                                "4096-request batch; never part of value, e2e or the kernel rooflines"}
        except Exception as exc:  # noqa: BLE001
            api_edge = {"error": repr(exc)[:200]}

    # ---- every kernel of the path alone, at sizes where bandwidth (not launch latency) is the limit -------------
    kernels = None
    if world == 1 and not args.no_kernels:
        del kd, xd, out_ring, hit_ring
        runner = pipe = None
        torch.cuda.empty_cache()
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import bench_kernels
        kernels = bench_kernels.run()

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline_single()

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "requests/s", "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "tf32" if MODEL == "linear_tf32" else "f32",
                "data": "synthetic", "config": dict(workload_config(world), hit_rate=round(hit_rate, 4), model=MODEL),
                "gpu_launches": int(launches), "host_enqueue_us_per_step": round(1e6 * t_host / K, 2),
                "clocks": clocks, "e2e": e2e, "roofline": roofline, "stages": stages,
                "kernels": kernels, "api_edge": api_edge, "cpu_baseline": cpu}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
