import os, sys
os.environ["DIE_HOST_TRACE"] = "1"
ROOT = os.getcwd(); sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-inference-engine_b200")]
import numpy as np, torch, bench
from die.pipeline import DispatchPipeline
B, D, G = bench.B, bench.D, 128
w, b = bench.model_weights()
pipe = DispatchPipeline(1 << 20, B, D, D, "linear_tf32", w, b, 300.0)
rng = np.random.default_rng(0)
_, keys, off, x = bench.make_batches(rng, 2 * G)
hk = torch.from_numpy(np.concatenate([keys, np.zeros((2 * G, 16), np.uint8)], axis=1)).pin_memory()
ho = torch.from_numpy(off).pin_memory(); hx = torch.from_numpy(x).pin_memory()
hout = torch.empty((G, B, D), dtype=torch.float32).pin_memory(); hhit = torch.empty((G, B), dtype=torch.uint8).pin_memory()
os.environ["DIE_HOST_TRACE"] = "1"
import time
for it in range(3):
    sys.stderr.write("call %d\n" % it)
    t = time.perf_counter()
    pipe.run_host_many(hk[(it & 1) * G:(it & 1) * G + G], ho, hx[(it & 1) * G:(it & 1) * G + G], 1.0 + it, 0.001, hout, hhit)
    sys.stderr.write("wall %.3f ms\n" % ((time.perf_counter() - t) * 1e3))
