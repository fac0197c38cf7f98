"""Diagnostic (not part of the product): phase times of the persistent step kernel at the C2 configuration, from
the %globaltimer marks CTA 0 leaves for the last batch of a launch."""
import ctypes, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-inference-engine_b200")]
import numpy as np, torch
import bench
from die import _cabi
from die.pipeline import DispatchPipeline

B, D = bench.B, bench.D
w, b = bench.model_weights()
kvmax = int(os.environ.get("DIAG_KV", bench.KV_MAX))
pipe = DispatchPipeline(kvmax, B, D, D, "linear_tf32", w, b, 300.0)
lib = _cabi.load()
rng = np.random.default_rng(0)
off_d = torch.from_numpy(np.arange(B + 1, dtype=np.int64) * 15).cuda()
out = torch.empty((64, B, D), device="cuda"); hit = torch.empty((64, B), dtype=torch.uint8, device="cuda")
pool = torch.from_numpy(rng.standard_normal((128, B, D), dtype=np.float32)).cuda()
now = 0.0
n_fill = int(kvmax * 1.15 / B) + 1
_, keys, _, _ = bench.make_batches(rng, n_fill + 128 * 6, payload=False)
kd = torch.from_numpy(keys).cuda()
for c in range(0, n_fill, 128):
    m = min(128, n_fill - c)
    pipe.run_device_many(kd[c:c + m], off_d, pool[:m], now, 0.01, out, hit); now += m * 0.01
torch.cuda.synchronize()
marks = (ctypes.c_uint64 * 8)()
for rep in range(6):
    lo = n_fill + rep * 128
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    th = time.perf_counter()
    pipe.run_device_many(kd[lo:lo + 128], off_d, pool, now, 0.01, out, hit); now += 1.28
    th = time.perf_counter() - th
    e1.record(); torch.cuda.synchronize()
    print("host enqueue %.1f us for 128 batches |" % (1e6 * th), end=" ")
    lib.die_pipeline_debug_marks_sync(pipe._h, marks)
    m = list(marks)
    us = lambda a, b_: (m[b_] - m[a]) / 1e3
    print("%.2f us/batch | last batch: probe issue %.2f  wait B1 %.2f | insert %.2f  wait B2 %.2f | epilogue %.2f  wait B3 %.2f | total %.2f" % (
        1e3 * e0.elapsed_time(e1) / 128, us(0, 4), us(4, 1), us(1, 5), us(5, 2), us(2, 6), us(6, 3), us(0, 3)))
NC = 144
cm = (ctypes.c_uint64 * (NC * 8))()
lib.die_pipeline_debug_cta_marks_sync(pipe._h, NC, cm)
t = np.array(list(cm), dtype=np.float64).reshape(NC, 8)
t0 = t[:, 0].min()
names = {0: "start", 4: "probes issued", 1: "B1 passed", 5: "inserts done", 7: "acc ready / bitmap", 2: "B2 passed", 6: "epilogue done", 3: "B3 passed"}
for role, sl in (("request CTAs", slice(0, 128)), ("window CTAs", slice(128, 144))):
    print(role)
    for i in (0, 4, 1, 5, 7, 2, 6, 3):
        v = (t[sl, i] - t0) / 1e3
        print("  %-20s min %6.2f  med %6.2f  max %6.2f us" % (names[i], v.min(), np.median(v), v.max()))
print(pipe.kv.stats())
