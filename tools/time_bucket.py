"""Diagnostic: time die_bucket_by_shard and die_route_md5 at the exchange's batch size."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-inference-engine_b200")]
import numpy as np, torch
import bench
from die import ops
rng = np.random.default_rng(0)


def timed(fn, n=100):
    """GPU time per call with the host out of the way: n calls captured into one graph, replayed."""
    for _ in range(3): fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    side = torch.cuda.Stream()
    with torch.cuda.stream(side):
        with torch.cuda.graph(g, stream=side):
            for _ in range(n): fn()
    g.replay(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); g.replay(); e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / n


B = 4096
for n in (2, 8, 16, 20, 40):
    shard = torch.from_numpy(rng.integers(0, n, B).astype(np.int32)).cuda()
    print("bucket n_shards", n, "us/launch %.2f" % timed(lambda: ops.bucket_by_shard(shard, n)))
_, keys, _, _ = bench.make_batches(rng, 1)
kb = ops.KeyBatch(torch.from_numpy(keys[0]).cuda(), torch.from_numpy(np.arange(B + 1, dtype=np.int64) * 15).cuda())
print("route us/launch %.2f" % timed(lambda: ops.route_md5(kb, 8)))
