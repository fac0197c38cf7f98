"""Diagnostic: device-side start times of the kernels of graph-replayed steps (die_debug_trace_*).
torchrun for N > 1 (sharded step), plain python for the single-GPU pipeline."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-inference-engine_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import bench  # noqa: E402
from die import _cabi  # noqa: E402

NAMES = {1: "route", 2: "dispatch", 3: "bar_in", 4: "bar_out", 5: "get", 6: "get_tail", 7: "claim", 8: "model",
         9: "unpack", 10: "scatter", 11: "step_end"}
rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(rank)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
B, D = bench.B, bench.D
w, b = bench.model_weights()
if world > 1:
    from die.exchange import PeerDispatcher
    runner = PeerDispatcher(world, rank, 1 << 20, B, D, w, b, 300.0, model="linear_tf32")
    run = runner.step_many
else:
    from die.pipeline import DispatchPipeline
    runner = DispatchPipeline(1 << 20, B, D, D, "linear_tf32", w, b, 300.0)
    run = runner.run_device_many
rng = np.random.default_rng(rank)
off_d = torch.from_numpy(np.arange(B + 1, dtype=np.int64) * 15).cuda()
NW, NT = int(os.environ.get('NW', '480')), int(os.environ.get('NT', '64'))
_, keys, _, x = bench.make_batches(rng, NW + NT)
kd, xd = torch.from_numpy(keys).cuda(), torch.from_numpy(x).cuda()
out = torch.empty((64, B, D), device="cuda")
hit = torch.empty((64, B), dtype=torch.uint8, device="cuda")
for c in range(0, NW, 64):
    run(kd[c:c + 64], off_d, xd[c:c + 64], 1.0 + c * 0.01, 0.01, out, hit)
torch.cuda.synchronize()
lib = _cabi.load()
_cabi.check(lib.die_debug_trace_enable(8192))
if world > 1:
    dist.barrier()
torch.cuda.synchronize()
run(kd[NW:], off_d, xd[NW:], 1.0 + NW * 0.01, 0.01, out, hit)
torch.cuda.synchronize()
print(rank, runner.pipe.kv.stats() if world > 1 else runner.kv.stats())
cap = 8192
tag, ns, n = (ctypes.c_uint64 * cap)(), (ctypes.c_uint64 * cap)(), ctypes.c_uint64()
_cabi.check(lib.die_debug_trace_read_sync(cap, tag, ns, ctypes.byref(n)))
if rank == 0:
    recs = sorted((ns[i], tag[i] >> 8, tag[i] & 0xff) for i in range(n.value))
    t0 = recs[0][0]
    print("records", len(recs), "span us", (recs[-1][0] - t0) / 1e3, "per step", (recs[-1][0] - t0) / 1e3 / NT)
    # mean interval from each kind of mark to the next mark on the device
    import collections
    acc = collections.defaultdict(list)
    for (ta, ia, ka), (tb, ib, kb) in zip(recs, recs[1:]):
        acc[(NAMES.get(ka, ka), NAMES.get(kb, kb))].append((tb - ta) / 1e3)
    if world == 1:
        inner = [(tb - ta) / 1e3 for (ta, ia, ka), (tb, ib, kb) in zip(recs, recs[1:]) if ka == 8 and kb == 5 and ib != 0]
        edge = [(tb - ta) / 1e3 for (ta, ia, ka), (tb, ib, kb) in zip(recs, recs[1:]) if ka == 8 and kb == 5 and ib == 0]
        print("model->get inside a replay: mean %.1f max %.1f | across replays: mean %.1f median %.1f" % (
            sum(inner) / len(inner), max(inner), sum(edge) / len(edge), sorted(edge)[len(edge) // 2]))
        print("across replays, sorted:", " ".join("%.0f" % v for v in sorted(edge)))
        for k2, v in sorted(acc.items()):
            print(k2, "n", len(v), "mean us %.1f" % (sum(v) / len(v)), "max %.1f" % max(v))
    if os.environ.get("ONLY"):  # e.g. ONLY=5,11: every owner launch's start and end over the whole traced queue
        kinds = {int(v) for v in os.environ["ONLY"].split(",")}
        for t, i, k in recs:
            if k in kinds:
                print(f"{(t - t0) / 1e3:9.1f}  {i & 0xffff:5d} {NAMES.get(k, k)}")
        sys.exit(0) if world == 1 else None
    lo = recs[len(recs) // 2][0]
    for t, i, k in recs:
        if lo <= t < lo + int(os.environ.get('WIN', '450')) * 1000:
            step = "ep%-6d" % i if k in (3, 4) else "step%-4d" % i
            print(f"{(t - lo) / 1e3:8.1f}  {step} {NAMES.get(k, k)}")
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
