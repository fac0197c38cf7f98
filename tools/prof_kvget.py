"""Profiling driver: batched get on a large table (run under ncu -k regex:kv_get_kernel)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-inference-engine_b200")]
import numpy as np, torch
import bench
from die import ops
from die.kvstore import DeviceKV
n_entries = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
B = int(sys.argv[2]) if len(sys.argv) > 2 else 1 << 20
rng = np.random.default_rng(0)
def dig_of(ids):
    kb = torch.from_numpy(bench.key_bytes_of(ids).reshape(-1)).cuda()
    off = torch.arange(ids.size + 1, dtype=torch.int64, device="cuda") * 15
    return ops.route_md5(ops.KeyBatch(kb, off), 1)[1]
kv = DeviceKV(n_entries, 300.0)
for a in range(0, n_entries, 1 << 20):
    m = min(1 << 20, n_entries - a)
    kv.set_batch(dig_of(np.arange(a, a + m, dtype=np.int64)), torch.arange(a, a + m, dtype=torch.int64, device="cuda"), 1.0)
ids = np.concatenate([rng.integers(0, n_entries, int(B * 0.9)), rng.integers(n_entries, 2 * n_entries, B - int(B * 0.9))])
rng.shuffle(ids)
d = dig_of(ids)
for i in range(4):
    kv.get_batch(d, 2.0 + i)
torch.cuda.synchronize()
# random-gather ceiling: 64-byte rows gathered from a 4 GB buffer by a library kernel (calibration only)
big = torch.empty((1 << 26, 16), dtype=torch.float32, device="cuda")
idx = torch.randint(0, 1 << 26, (B,), device="cuda")
for _ in range(3):
    g = big[idx]
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    g = big[idx]
e1.record(); torch.cuda.synchronize()
t = e0.elapsed_time(e1) / 5 * 1e-3
print(f"library random 64-byte gather: {B} rows in {t*1e6:.1f} us = {B/t/1e9:.2f} G rows/s, {B*64/t/1e9:.0f} GB/s of payload")
print(kv.stats())
