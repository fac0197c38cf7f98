"""Diagnostic (not part of the product): where does a C2 step spend its time -- host enqueue or device?"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-inference-engine_b200")]
import numpy as np, torch
import bench
from die import ops, _cabi
from die.pipeline import DispatchPipeline

B, D = bench.B, bench.D
w, b = bench.model_weights()
kvmax = int(sys.argv[1]) if len(sys.argv) > 1 else bench.KV_MAX
pipe = DispatchPipeline(kvmax, B, D, D, bench.MODEL, w, b, 300.0)
rng = np.random.default_rng(0)
off_d = torch.from_numpy(np.arange(B + 1, dtype=np.int64) * 15).cuda()
out = torch.empty((B, D), device="cuda"); hit = torch.empty(B, dtype=torch.uint8, device="cuda")
now = 0.0
n_fill = int(kvmax * 1.15 / B) + 1
for c in range(0, n_fill, 32):
    _, keys, _, x = bench.make_batches(rng, min(32, n_fill - c))
    kd, xd = torch.from_numpy(keys).cuda(), torch.from_numpy(x).cuda()
    for i in range(kd.shape[0]):
        now += 0.01; pipe.run_device(ops.KeyBatch(kd[i], off_d), xd[i], now, out, hit)
torch.cuda.synchronize()
N = 200
_, keys, _, x = bench.make_batches(rng, N)
kd, xd = torch.from_numpy(keys).cuda(), torch.from_numpy(x).cuda()
kbs = [ops.KeyBatch(kd[i], off_d) for i in range(N)]
xs = [xd[i] for i in range(N)]
torch.cuda.synchronize()
t0 = time.perf_counter()
for i in range(N):
    now += 0.01; pipe.run_device(kbs[i], xs[i], now, out, hit)
t1 = time.perf_counter()
torch.cuda.synchronize()
t2 = time.perf_counter()
print(f"enqueue {1e6*(t1-t0)/N:.1f} us/step, total {1e6*(t2-t0)/N:.1f} us/step")
# raw ctypes cost: route only
lib = _cabi.load()
sh = torch.empty(B, dtype=torch.int32, device="cuda"); dg = torch.empty((B, 4), dtype=torch.int32, device="cuda")
s = torch.cuda.current_stream().cuda_stream
torch.cuda.synchronize(); t0 = time.perf_counter()
for i in range(N):
    lib.die_route_md5(kd[i].data_ptr(), off_d.data_ptr(), B, 1, None, None, 0, 0, sh.data_ptr(), dg.data_ptr(), None, s)
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print(f"route only: enqueue {1e6*(t1-t0)/N:.1f} us, total {1e6*(t2-t0)/N:.1f} us")
# torch elementwise kernel back-to-back for reference
a = torch.zeros(1024, device="cuda")
torch.cuda.synchronize(); t0 = time.perf_counter()
for i in range(N * 9):
    a.add_(1.0)
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print(f"torch add_: enqueue {1e6*(t1-t0)/(N*9):.1f} us, total {1e6*(t2-t0)/(N*9):.1f} us per launch")
import ctypes
marks = (ctypes.c_uint64 * 8)()
_, keys3, _, x3 = bench.make_batches(rng, 6)
kd3, xd3 = torch.from_numpy(keys3).cuda(), torch.from_numpy(x3).cuda()
for rep in range(6):
    now += 0.01; pipe.run_device(ops.KeyBatch(kd3[rep], off_d), xd3[rep], now, out, hit)
    lib.die_pipeline_debug_marks_sync(pipe._h, marks)
    m = list(marks)
    print("get kernel marks (us): hash %.1f probe %.1f scan %.1f window %.1f (victims %d)" % tuple(
        [(m[i + 1] - m[i]) / 1e3 for i in range(4)] + [m[5]]))
print(pipe.kv.stats())
