"""Profiling driver (not part of the product): the C3-shaped ragged pack / unpack launches, for ncu."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-inference-engine_b200")]
import numpy as np, torch
from die import ops

rng = np.random.default_rng(3)
n_req = 64 * 256
data = []
for r in range(2):
    lens = rng.integers(16, 2049, n_req)
    off = np.zeros(n_req + 1, dtype=np.int64)
    np.cumsum(lens, out=off[1:])
    vals = torch.empty(int(off[-1]) + 8, dtype=torch.float32, device="cuda").normal_()
    data.append((vals, torch.from_numpy(off).cuda()))
flat = [torch.empty_like(d[0]) for d in data]
for i in range(6):
    v, o = data[i % 2]
    packed, _ = ops.pack_ragged(v, o, 2048)
    ops.unpack_ragged(packed, o, flat[i % 2])
torch.cuda.synchronize()
assert torch.equal(flat[1][: int(data[1][1][-1])], data[1][0][: int(data[1][1][-1])])
print("ok")
