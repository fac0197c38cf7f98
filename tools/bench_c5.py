"""BASELINE.json configs[4] as a path: the 4096-hidden bf16 MLP behind the sharded step, one hash shard per GPU.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P tools/bench_c5.py

Every rank ingests batches of 1024 requests (key + 4096-dim fp32 payload row), routes them by MD5 mod N, ships the rows
to their owners over NVLink, the owners probe their caches, run the misses through the tcgen05 MLP (bf16 storage, fp32
accumulation), cache and return the results.  Prints one JSON line (rank 0): requests/s, the model FLOP the misses needed
and the FLOP the GEMMs executed (they run over owner_max rows; the miss count stays on the device)."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-inference-engine_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from die.exchange import PeerDispatcher  # noqa: E402

rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
B, D, H = 1024, 4096, 4096
KV, OWNER_MAX = 1 << 15, 1280          # owners receive 1024 +- 32 requests per step
N_WARM, N_TIMED = 64, 128
rng = np.random.default_rng(0)
w1 = (rng.standard_normal((H, D)) / np.sqrt(D)).astype(np.float32)
w2 = (rng.standard_normal((D, H)) / np.sqrt(H)).astype(np.float32)
b1, b2 = rng.standard_normal(H).astype(np.float32), rng.standard_normal(D).astype(np.float32)
disp = PeerDispatcher(world, rank, KV, B, D, (w1, w2), (b1, b2), 300.0, model="mlp_bf16", owner_max=OWNER_MAX)
rng = np.random.default_rng(100 + rank)
n = N_WARM + N_TIMED
ids = rng.integers(0, 10 * KV * world, (n, B))          # ~10 % of a full table's worth of ids hit
keys = np.frombuffer("".join("user-%010d" % i for i in ids.ravel()).encode(), dtype=np.uint8).reshape(n, B * 15)
kd = torch.from_numpy(np.concatenate([keys, np.zeros((n, 16), np.uint8)], axis=1)).cuda()
off = torch.from_numpy(np.arange(B + 1, dtype=np.int64) * 15).cuda()
pool = torch.randn((16, B, D), device="cuda")             # payload rows: 16 batches, reused round-robin (256 MB)
out = torch.empty((16, B, D), device="cuda")
hit = torch.empty((16, B), dtype=torch.uint8, device="cuda")


def run(lo, hi, now):
    for c in range(lo, hi, 16):
        disp.step_many(kd[c:c + 16], off, pool, now + c * 0.01, 0.01, out, hit)


run(0, N_WARM, 1.0)
torch.cuda.synchronize()
s0 = disp.pipe.kv.stats()
dist.barrier()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
run(N_WARM, n, 1.0)
e1.record()
torch.cuda.synchronize()
dist.barrier()
disp.check()
s1 = disp.pipe.kv.stats()
t = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device="cuda")
dist.all_reduce(t, op=dist.ReduceOp.MAX)
cnt = torch.tensor([s1["misses"] - s0["misses"], s1["hits"] - s0["hits"]], dtype=torch.float64, device="cuda")
dist.all_reduce(cnt)
if rank == 0:
    sec = float(t[0]) / 1e3
    flop_row = 2.0 * 2.0 * D * H
    print(json.dumps({
        "config": "C5: 4096-hidden bf16 MLP behind the sharded step, %d requests per rank and step, %d GPUs" % (B, world),
        "n_gpus": world, "steps": N_TIMED, "ms_per_step": 1e3 * sec / N_TIMED,
        "requests_per_s": world * N_TIMED * B / sec, "hit_rate": float(cnt[1] / (cnt[0] + cnt[1])),
        "model_tflops_needed": float(cnt[0]) * flop_row / sec / 1e12,
        "model_tflops_executed": world * N_TIMED * OWNER_MAX * flop_row / sec / 1e12,
        "owner_max": OWNER_MAX, "dtype": "bf16 (fp32 accumulation), fp32 rows at the interface", "data": "synthetic"}))
dist.barrier()
dist.destroy_process_group()
