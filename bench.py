#!/usr/bin/env python
"""Headline benchmark: requests/sec of the batched route -> probe -> pack -> infer -> unpack path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (BASELINE.json configs[1], "C2"): flushed batches of 4096 requests, key "user-%010d" with ids
uniform in [0, 1e7), payload 128 fp32, mock linear model 128->128, one owner KV cache of 2^20 entries
per GPU (LRU, default_ttl 300 s), so about 10 % of the requests hit the cache and every batch evicts.
One STEP = a queue of G = 128 flushed batches (524 288 requests) handed to the path in one call; every batch
is still its own pass of the path, applied strictly in order (">= 200 back-to-back
batches").  Keys are freshly drawn for every batch; payload rows come from a pool of two steps' worth of
batches (537 MB, larger than L2) used alternately.  All inputs are resident in HBM before the timed region.

Prints ONE JSON line (see the contract in the task description): `value` is device-resident
throughput, `e2e` goes through the C-ABI with host buffers, `roofline` describes the dominant kernel
timed with CUDA events on its own stream, `cpu_baseline` is the reference's CPU path timed on this host.
"""
import argparse
import json
import os
import sys
import threading
import time

# The CPU arms are one Python process per core: a BLAS thread pool inside each of them oversubscribes the host
# (round 1: 8-12x slower reference numbers).  Must be set before numpy is imported.
if "--impl" in sys.argv and "reference" in sys.argv:
    for _v in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS", "NUMEXPR_NUM_THREADS"):
        os.environ[_v] = "1"

# The sharded step keeps one persistent owner kernel resident per replayed group while its ingest / return kernels run
# beside it on other streams: every stream needs its own hardware queue (must be set before CUDA is initialised).
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

ROOT = os.path.dirname(os.path.abspath(__file__))
for _p in (ROOT, os.path.join(ROOT, "distributed-inference-engine_b200")):
    if _p not in sys.path:
        sys.path.insert(0, _p)

import numpy as np  # noqa: E402

B, D = 4096, 128
G = int(os.environ.get("DIE_BENCH_BATCHES_PER_STEP", "128"))  # flushed batches per step (one queue call)
KV_MAX, ID_SPACE, TTL = 1 << 20, 10 ** 7, 300.0
KEY_LEN = 15
WORLD = int(os.environ.get("WORLD_SIZE", "1"))
METRIC = "requests/sec, batched route->probe->pack->infer->unpack"
MODEL = os.environ.get("DIE_BENCH_MODEL", "linear_tf32")  # tcgen05 kind::tf32 GEMM, fp32 accumulate
DT = 0.01  # seconds of injected clock per flushed batch


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
# CPU arms: the reference's pure-Python path on the host cores -- the one place bench.py may run oracle/
# ----------------------------------------------------------------------------------------------------------------
def cpu_kind():
    """"reference": the reference's own classes (verbatim copy under oracle/_ref, see oracle/fetch_ref.py);
    "port": the restatement in oracle/ref_port.py when that copy is absent."""
    from oracle import refrun
    return "reference" if refrun.available() else "port"


def cpu_worker(seed, n_timed):
    """One single-threaded replica of the path at the GPU arm's configuration: a 2^20-entry cache prefilled to
    capacity from ids in [0, 1e7) (same ~10 % hit rate, every batch evicts), then n_timed flushed batches of
    4096 requests timed.  Returns (requests, seconds, kind)."""
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(limits=1)   # numpy's BLAS stays on this process's one core
    except Exception:  # noqa: BLE001
        pass
    from oracle import refrun
    w, b = model_weights()

    def model(rows):
        if not rows:
            return []
        y = (np.stack(rows) @ w.T + b).astype(np.float32)
        return [y[i] for i in range(len(rows))]

    kind = cpu_kind()
    if kind == "reference":
        pipe = refrun.ReferencePipeline(1, KV_MAX, TTL, model, max_batch_size=B)
    else:
        from oracle.pipeline import BaselinePipeline
        pipe = BaselinePipeline(1, KV_MAX, TTL, model, max_batch_size=B)
    rng = np.random.default_rng(1000 + seed)
    n_fill = int(KV_MAX * 1.15 / B) + 1
    now = 1000.0
    # prefill without the model (untimed): the cached value's content does not matter to the timed region
    fill_row = np.zeros(D, dtype=np.float32)
    pipe_model, pipe.model = pipe.model, (lambda rows: [fill_row] * len(rows))
    for _ in range(n_fill):
        keys = ["user-%010d" % i for i in rng.integers(0, ID_SPACE, B)]
        now += DT
        pipe.process(keys, [fill_row] * B, now)
    pipe.model = pipe_model
    batches = []
    for _ in range(n_timed):
        ids = rng.integers(0, ID_SPACE, B)
        batches.append((["user-%010d" % i for i in ids], list(rng.standard_normal((B, D), dtype=np.float32))))
    t0 = time.perf_counter()
    for keys, rows in batches:
        now += DT
        pipe.process(keys, rows, now)
    dt = time.perf_counter() - t0
    return n_timed * B, dt, kind


def cpu_stage_times(n_rep=3):
    """Per-stage time of the reference's path on one host core (oracle port), one 4096-request batch each:
    the CPU column beside the per-kernel GPU numbers."""
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
    return best


def _time_once(fn):
    t0 = time.perf_counter()
    fn()
    return time.perf_counter() - t0


def cpu_baseline_single(n_timed=24):
    n, dt, kind = cpu_worker(0, n_timed)
    per_batch = dt / n_timed
    try:
        stages = cpu_stage_times()
    except Exception as exc:  # noqa: BLE001
        stages = {"error": repr(exc)[:200]}
    return {"value": n / dt, "unit": "requests/s", "cores": 1, "kind": kind,
            "as_shipped": {"value": B / (per_batch + 0.1), "unit": "requests/s",
                           "note": "the same path with the reference's own sleep left in: mock_batch_inference waits "
                                   "100 ms per batch (mock_inference.py:12-73), added analytically to the measured batch time"},
            "stages": stages,
            "sample": f"{n_timed} batches x {B} requests after prefilling the 2^20-entry cache (ids in [0,1e7), the GPU "
                      f"arm's configuration); {_cpu_path_text(kind)}, sleeps removed, one thread"}


def _cpu_path_text(kind):
    if kind == "reference":
        return ("the reference's own Router + KVCache + Batcher classes (oracle/_ref, verbatim copy) composed by "
                "oracle/refrun.py, numpy fp32 linear as the batch callback")
    return ("oracle/pipeline.py BaselinePipeline (restatement: hashlib.md5 + dict/OrderedDict + asyncio batcher + numpy "
            "fp32 linear)")


def usable_cores():
    try:
        return len(os.sched_getaffinity(0))
    except Exception:  # noqa: BLE001
        return os.cpu_count() or 1


def run_reference(args):
    """`--impl reference`: the reference's pure-Python path on all usable host cores (at most 32), one
    single-threaded one-shard replica process per core, each at the GPU arm's per-shard configuration."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import multiprocessing as mp
    cores = max(1, min(usable_cores(), 32))
    n_timed = 24          # a bounded sample: 24 batches x 4096 requests per replica, whatever --steps says
    t0 = time.perf_counter()
    with mp.get_context("fork").Pool(cores) as pool:
        res = pool.starmap(cpu_worker, [(s, n_timed) for s in range(cores)])
    total = sum(r[0] for r in res)
    slowest = max(r[1] for r in res)
    kind = res[0][2]
    value = total / slowest
    line = {"metric": METRIC, "value": value, "unit": "requests/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * G * B / value, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic", "impl": "reference",
            "config": workload_config(args.gpus),
            "cpu_baseline": {"value": value, "unit": "requests/s", "cores": cores, "kind": kind,
                             "sample": f"{cores} single-threaded replica processes x {n_timed} batches x {B} requests, each "
                                       f"after prefilling its own 2^20-entry cache; {_cpu_path_text(kind)}; ms_per_step = "
                                       f"time for one step's {G * B} requests at the aggregate rate; "
                                       f"wall {time.perf_counter() - t0:.1f}s"},
            "e2e": {"value": value, "unit": "requests/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))


def workload_config(n_gpus):
    return {"workload": "C2: 4096-request flushed batches per GPU, key 'user-%010d' (ids uniform in [0,1e7 x n_gpus)), "
                        "128-dim fp32 payload, mock linear 128->128, LRU+TTL cache of 2^20 entries per GPU "
                        "(~10% hits, evicting)",
            "batch": B, "dim": D, "kv_max_size": KV_MAX, "id_space": ID_SPACE * n_gpus, "default_ttl": TTL,
            "model": "linear 128->128 (GPU arm: tf32 tensor-core inputs, fp32 accumulate; CPU arm: numpy fp32)",
            "batches_per_step": G, "requests_per_step": G * B,
            "parallelism": f"hash-shard x{n_gpus}",
            "l2_policy": "inputs larger than L2: fresh keys every batch, payload pool of 2 steps (537 MB) alternated"}


# ----------------------------------------------------------------------------------------------------------------
# clocks
# ----------------------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """SM clock and throttle reasons while the timed region executes (pynvml = what nvidia-smi reads), every
    `period` (12 ms: a 20-step region of ~55 ms gets 3-4 samples; the recipe's 200 ms would give none).

    An NVML query takes 2-3 ms during which a kernel launch issued by this process waits.  The timed region is
    a handful of long launches (one persistent kernel per 16 batches, ~0.3 ms each) that the host enqueues far ahead
    of the GPU, so a stalled launch call does not reach the device-side clock; bench.py prints the measured cost of the
    sampler (`nvml_ms_per_sample`) and DIE_BENCH_SAMPLER_MS=0 turns it off for an A/B."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.samples, self.reasons, self.stop_flag, self.max_mhz = [], set(), False, None
        self.period = float(os.environ.get("DIE_BENCH_SAMPLER_MS", "12")) / 1e3
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
        """One more sample right now (called when the last step of the timed region has been enqueued)."""
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
                "when": "every period_ms from the start of the timed region to its end, plus once when its last step was "
                        "enqueued"}


def bind_to_gpu_numa_node(local_rank):
    """Pins this process to the CPUs of the NUMA node its GPU hangs off (sysfs), so that the pinned host buffers it
    allocates afterwards are local to that GPU's PCIe root.  Returns the node, or None when the box does not say."""
    try:
        import glob
        import subprocess
        bus = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", str(local_rank)],
                             capture_output=True, text=True, timeout=20).stdout.strip().lower()
        if bus.startswith("00000000:"):
            bus = bus[4:]
        node = int(open("/sys/bus/pci/devices/%s/numa_node" % bus).read())
        if node < 0:
            return None
        cpus = set()
        for part in open("/sys/devices/system/node/node%d/cpulist" % node).read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return None
        os.sched_setaffinity(0, cpus)
        return node
    except Exception:  # noqa: BLE001
        return None


def ncu_traffic(kernel_name):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel_name`, from the committed
    `ncu --set full` captures of this bench command (profiles/ncu_full_r02_*.csv; the round-1 capture covers the
    three-kernel form).  The persistent kernel's launches are groups of 16 batches, in the capture as in the bench."""
    import csv
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    for fn in ("ncu_full_r02_bench.csv", "ncu_full_r02_pipe_step_kernel.csv", "ncu_full_r01_pipeline_kernels.csv"):
        path = os.path.join(ROOT, "profiles", fn)
        try:
            rows = list(csv.reader(open(path)))
            hdr, units = rows[0], rows[1]
            i_name, i_r, i_w = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
            vals = [float(r[i_r]) * scale[units[i_r]] + float(r[i_w]) * scale[units[i_w]]
                    for r in rows[2:] if len(r) == len(hdr) and kernel_name in r[i_name]]
            if vals:
                return round(sum(vals) / len(vals)), "profiles/%s (mean of %d launches)" % (fn, len(vals))
        except Exception:  # noqa: BLE001
            continue
    return None, None


def stage_rooflines(lib, ms_arr, cnt_arr, n_stages, n_prof, s0, s1, step_ms_per_batch):
    """Per-stage device time of n_prof stream-launched batches (CUDA events on the launching stream, see
    die_profile_begin) -> (stages, roofline of the dominant kernel).  Algorithmic bytes per launch = SURVEY.md
    section 8(d)'s per-request figures x the requests one launch processes."""
    hits = (s1["hits"] - s0["hits"]) / n_prof
    miss = (s1["misses"] - s0["misses"]) / n_prof
    evic = (s1["evictions"] - s0["evictions"]) / n_prof
    row = D * 4
    route_b = B * (KEY_LEN + 8 + 16)
    probe_b = hits * 121 + miss * 89
    alg = {
        "kv_get": route_b + probe_b + B * 5 + evic * 12,                                  # route + probe + rank + window
        "kv_set_claim": miss * 160 + B * 2 * row + B * 13,                                # insert + pack / hit copy
        "model": miss * (2 * row + 16),
        "unpack_store": miss * 3 * row + B * 9 + evic * 128,                              # unpack + cached row + evict
        # the fused step kernel: everything above once -- keys in, probe, insert (+ eviction), payload row in,
        # result row out, cached row out (misses) / cached row in (hits), weights amortised over the batch
        "fused_step": route_b + probe_b + miss * 160 + evic * 128 + miss * (3 * row) + hits * (2 * row) + 64 * 1024 / 16,
    }
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:  # noqa: BLE001
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    total_ms = sum(ms_arr[i] for i in range(n_stages))
    kernel_of = {"kv_get": "pipe_get_kernel", "kv_set_claim": "pipe_claim_pack_kernel", "model": "linear_tf32_kernel",
                 "unpack_store": "pipe_unpack_evict_kernel", "fused_step": "pipe_step_kernel"}
    stages = {}
    for i in range(n_stages):
        if cnt_arr[i] == 0:
            continue
        name = lib.die_profile_stage_name(i).decode()
        per = ms_arr[i] / cnt_arr[i]
        batches_per_launch = n_prof / cnt_arr[i] if name == "fused_step" else 1.0  # the persistent kernel: a group per launch
        alg_launch = alg.get(name, 0) * batches_per_launch
        gbs = alg_launch / (per * 1e-3) / 1e9 if per > 0 else 0.0
        stages[name] = {"kernel": kernel_of.get(name, name), "us_per_launch": round(per * 1e3, 3),
                        "launches_per_batch": cnt_arr[i] / n_prof, "batches_per_launch": round(batches_per_launch, 2),
                        "share": round(ms_arr[i] / total_ms, 4),
                        "alg_bytes_per_launch": round(alg_launch), "gbs": round(gbs, 1),
                        "frac_of_hbm_peak": round(gbs / peak, 4)}
    dom = max(stages, key=lambda n: stages[n]["share"])
    traffic, traffic_src = ncu_traffic(stages[dom]["kernel"])
    fused_alg = B * (KEY_LEN + 8 + 4) + probe_b + 2 * B * row     # SURVEY 8(d) fused C2 lower bound: 1.17 KB / request
    roofline = {"kernel": stages[dom]["kernel"], "bound": "hbm", "achieved": stages[dom]["gbs"], "peak": peak,
                "unit": "GB/s", "frac": stages[dom]["frac_of_hbm_peak"], "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": peak_src,
                "us_per_launch": stages[dom]["us_per_launch"], "alg_bytes_per_launch": stages[dom]["alg_bytes_per_launch"],
                "whole_step_fused_accounting": {
                    "alg_bytes_per_batch": round(fused_alg), "us_per_batch": round(step_ms_per_batch * 1e3, 3),
                    "gbs": round(fused_alg / (step_ms_per_batch * 1e-3) / 1e9, 1),
                    "frac": round(fused_alg / (step_ms_per_batch * 1e-3) / 1e9 / peak, 4),
                    "note": "SURVEY 8(d) north-star denominator (payload read once, result written once, route + probe) "
                            "over the device-timed batch of `value`"},
                "note": "a 4096-request batch carries ~5 MB of algorithmic traffic (<1 us at peak): the step sits on its "
                        "chain of dependent DRAM round trips and grid-wide barriers, see DESIGN.md section 5"}
    return stages, roofline


def e2e_sharded(runner, rng, K, now, off_d, world):
    """End to end at N > 1: every rank's batches start in pinned host memory and its results end there.  Groups of
    64 batches: one H2D copy per array, one die_xchg_step_many call, one D2H copy per array;
    three device staging sets on three streams keep the copy engines on the neighbouring groups.  Wall clock
    between two barriers, max over ranks."""
    import torch
    import torch.distributed as dist
    # 64 batches per call: die_xchg_step_many pipelines its groups of 16 across launches (ingest | owner | way back), so a
    # call of one group would be three serial launches (measured 1.2-1.4 ms per 16 batches beside the copies, more than
    # their PCIe time: 59 M requests/s at N = 2) while four groups per call cost six (101 M).  The warm-up is long enough
    # (24 calls) to take the tables' first housekeeping -- a one-time stall of ~0.5 s -- out of the timed region.
    GS = int(os.environ.get("DIE_E2E_GROUP", "64"))
    n_groups = max(2, K * G // GS)
    n = n_groups * GS
    n_pool = 2 * G                                    # payload pool (537 MB pinned per rank), keys fresh per batch
    _, keys, _, _ = make_batches(rng, n, payload=False)
    hk = torch.from_numpy(keys).pin_memory()
    hx = torch.from_numpy(rng.standard_normal((n_pool, B, D), dtype=np.float32)).pin_memory()
    hout = torch.empty((n_pool, B, D), dtype=torch.float32).pin_memory()
    hhit = torch.empty((n, B), dtype=torch.uint8).pin_memory()
    NS = 3
    dk = [torch.zeros((GS + 1, B * KEY_LEN), dtype=torch.uint8, device="cuda") for _ in range(NS)]  # +1 row of slack
    dx = [torch.empty((GS, B, D), dtype=torch.float32, device="cuda") for _ in range(NS)]
    dout = [torch.empty((GS, B, D), dtype=torch.float32, device="cuda") for _ in range(NS)]
    dhit = [torch.empty((GS, B), dtype=torch.uint8, device="cuda") for _ in range(NS)]
    s_in, s_out, cur = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.current_stream()
    ev_in = [torch.cuda.Event() for _ in range(NS)]
    ev_done = [torch.cuda.Event() for _ in range(NS)]
    ev_out = [torch.cuda.Event() for _ in range(NS)]

    trace = os.environ.get("DIE_E2E_TRACE") == "1"   # diagnostics: per-group copy-in / steps / copy-out times, rank 0
    marks = []

    def mark(stream):
        if trace:
            e = torch.cuda.Event(enable_timing=True)
            e.record(stream)
            marks.append(e)

    def run(first_group, count, now0):
        del marks[:]
        mark(cur)
        for g in range(first_group, first_group + count):
            b = g % NS
            lo, hi = (g % n_groups) * GS, (g % n_groups) * GS + GS
            plo = (g * GS) % n_pool
            with torch.cuda.stream(s_in):
                if g - first_group >= NS:
                    s_in.wait_event(ev_done[b])
                mark(s_in)
                dk[b][:GS].copy_(hk[lo:hi], non_blocking=True)
                dx[b].copy_(hx[plo:plo + GS], non_blocking=True)
                ev_in[b].record(s_in)
                mark(s_in)
            cur.wait_event(ev_in[b])
            if g - first_group >= NS:
                cur.wait_event(ev_out[b])
            mark(cur)
            runner.step_many(dk[b][:GS], off_d, dx[b], now0 + (g - first_group) * GS * DT, DT, dout[b], dhit[b])
            ev_done[b].record(cur)
            mark(cur)
            with torch.cuda.stream(s_out):
                s_out.wait_event(ev_done[b])
                mark(s_out)
                hout[plo:plo + GS].copy_(dout[b], non_blocking=True)
                hhit[lo:hi].copy_(dhit[b], non_blocking=True)
                ev_out[b].record(s_out)
                mark(s_out)
        s_out.synchronize()
        cur.synchronize()
        if trace and dist.get_rank() == 0:
            for g in range(min(count, 12)):
                t = [marks[0].elapsed_time(marks[1 + 6 * g + k]) for k in range(6)]
                print("group %2d  in %.3f-%.3f  steps %.3f-%.3f  out %.3f-%.3f" % (g, *t), file=sys.stderr)

    n_warm = int(os.environ.get("DIE_E2E_WARM_GROUPS", "24"))
    run(0, n_warm, now)                    # warm the staging path and the graphs
    now += n_warm * GS * DT
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
            "h2d_bytes_per_step": G * (B * KEY_LEN + B * D * 4), "d2h_bytes_per_step": G * (B * D * 4 + B),
            "ms_per_step": 1e3 * dt / n * G, "steps": n / G,
            "api": "PeerDispatcher.step_many (die_xchg_step_many) on groups of %d batches staged from / to pinned host " % GS +
                   "buffers on every rank; copy-in, sharded steps and copy-out of consecutive groups overlap"}, now + n * DT


def sharded_self_check(runner, rank, world, off_d):
    """The N > 1 arm checks itself before anything is timed (there is no multi-GPU pytest on a one-GPU test box):
    one stream-launched sharded step and one graph-replayed group of 16 on the empty owner tables, every rank's
    out / hit / shard compared with oracle.pipeline.OraclePipeline.process_exchange -- hit flags and shard ids
    bit-exact, outputs within 2**-10 * sum|x w| + 1e-6, owner counters equal.  A mismatch fails the run."""
    import torch
    from die import ops
    from oracle import models
    from oracle.pipeline import OraclePipeline
    w, b = model_weights()
    rel = 2.0 ** -10 if MODEL == "linear_tf32" else 1e-5

    def fn(rows):
        if not rows:
            return []
        xs = np.stack(rows)
        y, bound = models.linear_f64(xs, w, b), models.linear_abs_bound(xs, w)
        return [np.stack([y[i], bound[i]]) for i in range(len(rows))]

    ref = OraclePipeline(world, KV_MAX, TTL, fn)
    rng = np.random.default_rng(99)          # same stream on every rank: each rank knows all ranks' batches
    n = 17
    ids = rng.integers(0, 30000 * world, (n, world, B), dtype=np.int64)
    x = rng.standard_normal((n, world, B, D), dtype=np.float32)
    keys_of = lambda j, r: ["user-%010d" % i for i in ids[j, r]]  # noqa: E731
    now = 500.0

    def compare(j, out, hit, shard):
        want = ref.process_exchange([(keys_of(j, r), [row for row in x[j, r]]) for r in range(world)], now + j * DT)[rank]
        if shard is not None and shard.tolist() != want["shard"]:
            raise AssertionError(f"self-check: shard ids differ from the oracle at batch {j} on rank {rank}")
        if hit.tolist() != want["hit"]:
            raise AssertionError(f"self-check: hit flags differ from the oracle at batch {j} on rank {rank}")
        wv = np.stack(want["out"])
        if not np.all(np.abs(out.cpu().numpy().astype(np.float64) - wv[:, 0]) <= rel * wv[:, 1] + 1e-6):
            raise AssertionError(f"self-check: outputs outside tolerance at batch {j} on rank {rank}")

    kd = torch.from_numpy(key_bytes_of(ids[:, rank].reshape(-1)).reshape(n, B * KEY_LEN)).cuda()
    xd = torch.from_numpy(np.ascontiguousarray(x[:, rank])).cuda()
    out, hit, shard = runner.step(ops.KeyBatch(kd[0], off_d), xd[0], now)
    torch.cuda.synchronize()
    compare(0, out, hit, shard if isinstance(shard, torch.Tensor) else None)
    outs = torch.empty((n - 1, B, D), dtype=torch.float32, device="cuda")
    hits = torch.empty((n - 1, B), dtype=torch.uint8, device="cuda")
    if hasattr(runner, "step_many"):
        runner.step_many(kd[1:], off_d, xd[1:], now + DT, DT, outs, hits)
    else:
        for j in range(1, n):
            runner.step(ops.KeyBatch(kd[j], off_d), xd[j], now + j * DT, outs[j - 1], hits[j - 1])
    torch.cuda.synchronize()
    for j in range(1, n):
        compare(j, outs[j - 1], hits[j - 1], None)
    if hasattr(runner, "check"):
        runner.check()
    st, mine = runner.pipe.kv.stats(), ref.state()[rank]
    got, want = (st["hits"], st["misses"], st["evictions"], st["size"]), (mine["hits"], mine["misses"], mine["evictions"], mine["size"])
    if got != want:
        raise AssertionError(f"self-check: owner counters {got} != oracle {want} on rank {rank}")
    return {"batches": n, "requests_per_rank": n * B, "hits": int(mine["hits"]), "vs": "oracle.pipeline.OraclePipeline.process_exchange",
            "result": "ok"}


# ----------------------------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------------------------
def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
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

    self_check = None
    if world > 1:
        self_check = sharded_self_check(runner, rank, world, off_d)
        dist.barrier()

    def run_steps(kd, xd, now0):
        """kd [n, B*15] uint8, xd [n, B, D]: a queue of n flushed batches, applied in order."""
        n = kd.shape[0]
        if runner is None:  # one C-ABI call submits the whole queue (each batch is still its own pass)
            pipe.run_device_many(kd, off_d, xd, now0, DT, out_ring, hit_ring)
        elif hasattr(runner, "step_many"):
            runner.step_many(kd, off_d, xd, now0, DT, out_ring, hit_ring)
        else:
            for i in range(n):
                runner.step(ops.KeyBatch(kd[i], off_d), xd[i], now0 + i * DT, out_ring[i % RING], hit_ring[i % RING])
        return now0 + n * DT

    # ---- inputs: fresh keys for every batch; payload pool of two steps' worth of batches, larger than L2 ----------
    pool = torch.from_numpy(rng.standard_normal((2 * G, B, D), dtype=np.float32)).cuda()
    n_fill = int(KV_MAX * 1.15 / B) + 1
    _, keys, _, _ = make_batches(rng, n_fill + (W + K) * G, payload=False)
    kd_all = torch.from_numpy(keys).cuda()
    del keys

    # ---- prefill the cache to capacity (untimed) -------------------------------------------------------------
    now = 1000.0
    for c in range(0, n_fill, G):
        m = min(G, n_fill - c)
        now = run_steps(kd_all[c:c + m], pool[:m], now)
    torch.cuda.synchronize()
    kd_steps = kd_all[n_fill:].view(W + K, G, B * KEY_LEN)

    # ---- timed: inputs resident in HBM ------------------------------------------------------------------------
    for s_ in range(W):
        now = run_steps(kd_steps[s_], pool[(s_ & 1) * G:(s_ & 1) * G + G], now)
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
    for s_ in range(W, W + K):
        now = run_steps(kd_steps[s_], pool[(s_ & 1) * G:(s_ & 1) * G + G], now)
    t_host = time.perf_counter() - t_host  # host time to enqueue the queues (no wait on the GPU)
    ev1.record()
    sampler.sample_now()  # everything is enqueued: the queries can no longer hold up a launch
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
        runner.check() if hasattr(runner, "check") else None
    value = world * K * G * B / (ms / 1e3)
    hit_rate = (stats1["hits"] - stats0["hits"]) / max(1, (stats1["hits"] - stats0["hits"]) + (stats1["misses"] - stats0["misses"]))
    del kd_all, kd_steps

    # ---- e2e: host buffers through the C-ABI (H2D + kernels + D2H inside the timed region) ---------------------
    e2e = None
    if world > 1 and hasattr(runner, "step_many"):
        try:
            e2e, now = e2e_sharded(runner, rng, K, now, off_d, world)
        except Exception as exc:  # noqa: BLE001  (the device-resident number above stands on its own)
            e2e = {"value": None, "unit": "requests/s", "error": repr(exc)[:300]}
    if world == 1:
        _, keys, off, _ = make_batches(rng, (K + 1) * G, payload=False)
        hk = torch.from_numpy(np.concatenate([keys, np.zeros(((K + 1) * G, 16), np.uint8)], axis=1)).pin_memory()
        del keys
        ho = torch.from_numpy(off).pin_memory()
        hx = pool.cpu().pin_memory()                       # the payload pool, in pinned host memory
        hout = torch.empty((G, B, D), dtype=torch.float32).pin_memory()
        hhit = torch.empty((G, B), dtype=torch.uint8).pin_memory()
        pipe.run_host_many(hk[K * G:], ho, hx[:G], now, DT, hout, hhit)  # warm the staging path
        now += G * DT
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for s_ in range(K):    # one call per step: a queue of G batches, returns after the last result is on the host
            pipe.run_host_many(hk[s_ * G:(s_ + 1) * G], ho, hx[(s_ & 1) * G:(s_ & 1) * G + G], now, DT, hout, hhit)
            now += G * DT
        dt = time.perf_counter() - t0
        # the unpipelined form of the same call, one synchronous batch at a time
        t0 = time.perf_counter()
        n_sync = 50
        for i in range(n_sync):
            pipe.run_host(hk[i], ho, hx[i], now + i * DT, hout[i], hhit[i])
        dt_sync = time.perf_counter() - t0
        now += n_sync * DT
        e2e = {"value": K * G * B / dt, "unit": "requests/s",
               "h2d_bytes_per_step": G * (B * KEY_LEN + B * D * 4) + (B + 1) * 8, "d2h_bytes_per_step": G * (B * D * 4 + B),
               "ms_per_step": 1e3 * dt / K,
               "api": "die_pipeline_run_host_many (C-ABI, pinned host buffers), one call per step of G batches; copy-in / "
                      "kernels / copy-out of consecutive groups of %s batches overlap; each call returns with its results "
                      "on the host" % os.environ.get("DIE_HOST_GROUP", "16"),
               "per_call_sync": {"value": n_sync * B / dt_sync, "ms_per_batch": 1e3 * dt_sync / n_sync,
                                 "api": "die_pipeline_run_host, one blocking call per batch"}}
        del hk, hx, hout, hhit

    # ---- per-stage device time (CUDA events on the launching stream) and roofline -------------------------------
    stages, roofline = {}, None
    if world == 1 or os.environ.get("DIE_BENCH_PROFILE_ALL"):
        n_prof = 64   # four groups of 16 batches
        _, keys, _, _ = make_batches(rng, n_prof, payload=False)
        kd2, xd2 = torch.from_numpy(keys).cuda(), pool[:n_prof]
        s0 = pipe.kv.stats()
        _cabi.check(lib.die_profile_begin())
        now = run_steps(kd2, xd2, now)
        import ctypes
        ms_arr, cnt_arr, n_st = (ctypes.c_double * 32)(), (ctypes.c_uint64 * 32)(), ctypes.c_int32()
        _cabi.check(lib.die_profile_end(32, ms_arr, cnt_arr, ctypes.byref(n_st)))
        s1 = pipe.kv.stats()
        stages, roofline = stage_rooflines(lib, ms_arr, cnt_arr, n_st.value, n_prof, s0, s1, ms / (K * G))

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
        del out_ring, hit_ring, pool
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
                "ms_per_step": ms / K, "us_per_batch": round(1e3 * ms / (K * G), 3), "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None,
                "dtype": "tf32" if MODEL == "linear_tf32" else "f32",
                "data": "synthetic", "config": workload_config(world), "hit_rate": round(hit_rate, 4),
                "gpu_launches": int(launches), "host_enqueue_us_per_step": round(1e6 * t_host / K, 2),
                "clocks": clocks, "e2e": e2e, "roofline": roofline, "stages": stages,
                "kernels": kernels, "api_edge": api_edge, "self_check": self_check, "cpu_baseline": cpu}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
