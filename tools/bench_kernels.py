"""Per-kernel roofline measurements at bandwidth-relevant sizes (BASELINE.json configs[2]-[3] shapes).

Each stage is called through the same C-ABI entry points the product uses, timed with CUDA events on the
launching stream over several launches whose inputs/outputs rotate through buffers larger than L2
(126 MB), after warm-up.  achieved = algorithmic bytes per launch (SURVEY.md section 8(d) per-request
figures x requests per launch) / mean launch time; peak = MEASURED_PEAKS.json hbm_gbs.

    python tools/bench_kernels.py            # prints one JSON object
"""
import hashlib
import json
import math
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-inference-engine_b200")]

import numpy as np  # noqa: E402
import torch  # noqa: E402


def timed(fn, n_iter, warm=3):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n_iter):
        fn(warm + i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / n_iter  # seconds per launch


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:  # noqa: BLE001
        return 6650.0, "fallback 6650 GB/s (B200_PROFILING.md)"


def entry(name, secs, alg_bytes, units, unit_name, peak, note=""):
    gbs = alg_bytes / secs / 1e9
    return {"kernel": name, "us_per_launch": round(secs * 1e6, 2), "units_per_launch": units, "unit": unit_name,
            "alg_bytes_per_launch": int(alg_bytes), "gbs": round(gbs, 1), "frac_of_hbm_peak": round(gbs / peak, 4),
            "units_per_s": round(units / secs, 1), "note": note}


def run(kv_entries=10_000_000, quick=False):
    from die import _cabi, ops
    from die.kvstore import DeviceKV
    import bench
    peak, peak_src = peak_gbs()
    rng = np.random.default_rng(3)
    out = {"peak_gbs": peak, "peak_source": peak_src, "kernels": []}
    K = out["kernels"]

    # ---- echo (pure copy): calibrates what a streaming kernel of ours reaches -----------------------------------
    n = 64 << 20  # 256 MB in + 256 MB out
    xs = [torch.empty(n, dtype=torch.float32, device="cuda").normal_() for _ in range(2)]
    s = timed(lambda i: ops.echo_fwd(xs[i % 2]), 6)
    K.append(entry("echo_vec_kernel", s, 2 * n * 4, n * 4, "bytes", peak, "256 MB copy"))
    del xs

    # ---- random-sector ceiling over a 4 GB buffer (calibration kernel) -----------------------------------------
    lib = _cabi.require_cuda()
    big = torch.empty(1 << 30, dtype=torch.float32, device="cuda")  # 4 GB
    sink = torch.zeros(1, dtype=torch.int64, device="cuda")
    out["random_read_ceiling"] = {}
    for each in (32, 64):
        n_thr, per = 1 << 20, 16
        s = timed(lambda i: _cabi.check(lib.die_debug_random_read(big.data_ptr(), big.numel() * 4, n_thr, per, each, 1234 + i,
                                                                  sink.data_ptr(), _cabi.stream_ptr())), 5)
        acc = n_thr * per
        out["random_read_ceiling"][f"{each}B"] = {"accesses_per_s": round(acc / s, 1), "gbs": round(acc * each / s / 1e9, 1),
                                                  "frac_of_hbm_peak": round(acc * each / s / 1e9 / peak, 4)}
    del big

    # ---- route: 1M keys of 15 bytes ------------------------------------------------------------------------------
    nk = 1 << 20
    sets = []
    for r in range(3):
        ids = rng.integers(0, 10 ** 7, nk)
        kb = torch.from_numpy(bench.key_bytes_of(ids).reshape(-1)).cuda()
        off = torch.arange(nk + 1, dtype=torch.int64, device="cuda") * 15
        sets.append(ops.KeyBatch(kb, off))
    s = timed(lambda i: ops.route_md5(sets[i % 3], 8), 10)
    K.append(entry("route_md5_kernel", s, nk * (15 + 8 + 4 + 16), nk, "keys", peak,
                   "INT32-issue bound (about 700 integer ops per key); HBM fraction reported for completeness"))
    digests_1m = ops.route_md5(sets[0], 1)[1]

    # ---- pack / unpack, C3 shape: 64 flushes x 256 requests of 16..2048 fp32, Lpad 2048 ---------------------------
    n_req = 64 * 256
    data = []
    for r in range(2):
        lens = rng.integers(16, 2049, n_req)
        off = np.zeros(n_req + 1, dtype=np.int64)
        np.cumsum(lens, out=off[1:])
        vals = torch.empty(int(off[-1]) + 8, dtype=torch.float32, device="cuda").normal_()
        data.append((vals, torch.from_numpy(off).cuda(), int(off[-1])))
    packed = [None, None]

    def do_pack(i):
        v, o, _ = data[i % 2]
        packed[i % 2] = ops.pack_ragged(v, o, 2048)[0]
    s = timed(do_pack, 6)
    tot = sum(d[2] for d in data) / 2
    K.append(entry("pack_vec_kernel", s, 4 * tot + 8 * n_req + 4 * 2048 * n_req, n_req, "requests", peak,
                   "C3 ragged 16..2048 fp32 -> [16384, 2048]"))
    flat = [torch.empty_like(d[0]) for d in data]
    s = timed(lambda i: ops.unpack_ragged(packed[i % 2], data[i % 2][1], flat[i % 2]), 6)
    K.append(entry("unpack_vec_kernel", s, 2 * 4 * tot + 8 * n_req, n_req, "requests", peak,
                   "C3 padded rows -> ragged slices (reads only the live part of each row)"))
    del data, packed, flat

    # ---- pack, C2 shape: 1M fixed rows of 128 fp32 (gather by index) -----------------------------------------------
    n_req = 1 << 19
    vals = torch.empty(n_req * 128, dtype=torch.float32, device="cuda").normal_()
    off = torch.arange(n_req + 1, dtype=torch.int64, device="cuda") * 128
    perm = torch.randperm(n_req, device="cuda").to(torch.int32)
    s = timed(lambda i: ops.pack_ragged(vals, off, 128, perm), 6)
    K.append(entry("pack_vec_kernel (gather)", s, n_req * (512 + 8 + 4 + 512), n_req, "requests", peak,
                   "C2 rows of 512 B gathered through a random index"))
    del vals, off, perm

    # ---- linear tf32: 64k rows of 128 ------------------------------------------------------------------------------
    w, b = bench.model_weights()
    wd, bd = torch.from_numpy(w).cuda(), torch.from_numpy(b).cuda()
    xs = [torch.empty((1 << 17, 128), dtype=torch.float32, device="cuda").normal_() for _ in range(3)]
    s = timed(lambda i: ops.linear_fwd(xs[i % 3], wd, bd, _cabi.LINEAR_TF32), 8)
    rows = 1 << 17
    K.append(entry("linear_tf32_kernel", s, rows * 1024 + 65536, rows, "rows", peak, "131072 x 128 -> 128, tcgen05 kind::tf32, TMA in and out (linear_tf32_tma_kernel)"))
    del xs

    # ---- bf16 MLP (C5: hidden 4096, batch 8192): the tensor-bound kernel of the path --------------------------------
    try:
        tf_peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["bf16_tflops"])
        tf_src = "MEASURED_PEAKS.json bf16_tflops (cuBLAS 8192^3, burst)"
    except Exception:  # noqa: BLE001
        tf_peak, tf_src = 2250.0, "nominal dense bf16 2.25 PFLOP/s"
    Bm, Dm, Hm = 8192, 4096, 4096
    xs = [torch.empty((Bm, Dm), dtype=torch.bfloat16, device="cuda").normal_() for _ in range(2)]
    w1 = torch.empty((Hm, Dm), dtype=torch.bfloat16, device="cuda").normal_(0, Dm ** -0.5)
    w2 = torch.empty((Dm, Hm), dtype=torch.bfloat16, device="cuda").normal_(0, Hm ** -0.5)
    b1 = torch.zeros(Hm, dtype=torch.float32, device="cuda")
    b2 = torch.zeros(Dm, dtype=torch.float32, device="cuda")
    s = timed(lambda i: ops.mlp_fwd_bf16(xs[i % 2], w1, b1, w2, b2), 5)
    flops = 2 * 2.0 * Bm * Dm * Hm
    out["mlp_bf16"] = {"kernel": "gemm_bf16_kernel x2 (persistent, tcgen05 kind::f16, TMA, two TMEM accumulators)", "batch": Bm, "dim": Dm, "hidden": Hm,
                       "ms_per_forward": round(s * 1e3, 4), "tflops": round(flops / s / 1e12, 1), "peak_tflops": tf_peak,
                       "peak_source": tf_src, "frac_of_tensor_peak": round(flops / s / 1e12 / tf_peak, 4),
                       "rows_per_s": round(Bm / s, 1)}
    del xs, w1, w2

    # ---- KV probe / fill on a large table (C4: 10M entries, LRU + TTL, 90 % hits) -----------------------------------
    n_entries = kv_entries if not quick else 1_000_000
    kv = DeviceKV(n_entries, 300.0)
    chunk = 1 << 20
    now = 100.0
    filled = 0

    def digests_of_ids(ids):
        kb = torch.from_numpy(bench.key_bytes_of(ids).reshape(-1)).cuda()
        off = torch.arange(ids.size + 1, dtype=torch.int64, device="cuda") * 15
        return ops.route_md5(ops.KeyBatch(kb, off), 1)[1]

    while filled < n_entries:  # ids 0 .. n_entries-1 in order
        m = min(chunk, n_entries - filled)
        dig = digests_of_ids(np.arange(filled, filled + m, dtype=np.int64))
        kv.set_batch(dig, torch.arange(filled, filled + m, dtype=torch.int64, device="cuda"), now)
        filled += m
        now += 0.001
    assert kv.sync_status()[0] == 0
    for B in (65536, 1 << 20):
        batches = []
        for r in range(4):
            hit_ids = rng.integers(0, n_entries, int(B * 0.9))
            miss_ids = rng.integers(n_entries, 2 * n_entries, B - hit_ids.size)
            ids = np.concatenate([hit_ids, miss_ids])
            rng.shuffle(ids)
            batches.append(digests_of_ids(ids))
        st0 = kv.stats()
        tick = [now]

        def do_get(i):
            tick[0] += 0.001
            kv.get_batch(batches[i % 4], tick[0])
        n_it = 12
        s = timed(do_get, n_it)
        now = tick[0]
        st1 = kv.stats()
        hits = (st1["hits"] - st0["hits"]) / (n_it + 3)
        miss = (st1["misses"] - st0["misses"]) / (n_it + 3)
        K.append(entry("kv_get_kernel", s, hits * 121 + miss * 89, B, "probes", peak,
                       f"{n_entries} entries ({kv.stats()['capacity'] * 64 / 1e9:.1f} GB of slots), batch {B}, "
                       f"hit rate {hits / B:.3f}; random 32-byte sectors"))
    # fill: sets of absent keys into the full table -> every insert evicts the LRU head
    for M in (6554, 65536):
        fresh = [digests_of_ids(np.arange(3 * n_entries + r * M, 3 * n_entries + (r + 1) * M, dtype=np.int64) + M * 100 * (M % 7))
                 for r in range(10)]
        vals = torch.arange(M, dtype=torch.int64, device="cuda")
        tick = [now]

        def do_set(i):
            tick[0] += 0.001
            kv.set_batch(fresh[i % 10], vals, tick[0])
        s = timed(do_set, 6, warm=2)
        now = tick[0]
        K.append(entry("kv_set (precheck+window+claim+finalize+evict)", s, M * 288, M, "inserts", peak,
                       f"{M} new keys into the full table: {M} evictions; 7 launches (window and evictions over parallel tiles)"))
    assert kv.sync_status()[0] == 0
    st = kv.stats()
    out["kv_table"] = {"entries": st["size"], "capacity": st["capacity"], "evictions": st["evictions"]}
    return out


if __name__ == "__main__":
    print(json.dumps(run(quick="--quick" in sys.argv)))
