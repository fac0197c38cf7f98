"""Diagnostic: event timeline of the sharded step's launches on N GPUs (torchrun), stream-launched
(profiling turns the graph replays off).  Prints rank 0's launches of a few steady-state steps."""
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
from die.exchange import PeerDispatcher  # noqa: E402

rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
B, D = bench.B, bench.D
w, b = bench.model_weights()
disp = PeerDispatcher(world, rank, 1 << 18, B, D, w, b, 300.0, model="linear_tf32")
rng = np.random.default_rng(rank)
off_d = torch.from_numpy(np.arange(B + 1, dtype=np.int64) * 15).cuda()
N = 96
_, keys, _, x = bench.make_batches(rng, N)
kd, xd = torch.from_numpy(keys).cuda(), torch.from_numpy(x).cuda()
out = torch.empty((8, B, D), device="cuda")
hit = torch.empty((8, B), dtype=torch.uint8, device="cuda")
disp.step_many(kd[:64], off_d, xd[:64], 1.0, 0.01, out, hit)
torch.cuda.synchronize()
dist.barrier()
lib = _cabi.load()
_cabi.check(lib.die_profile_begin())
disp.step_many(kd[64:], off_d, xd[64:], 2.0, 0.01, out, hit)
cap = 4096
st, sm = (ctypes.c_int32 * cap)(), (ctypes.c_int32 * cap)()
t0, t1 = (ctypes.c_double * cap)(), (ctypes.c_double * cap)()
n = ctypes.c_int32()
_cabi.check(lib.die_profile_timeline(cap, st, sm, t0, t1, ctypes.byref(n)))
ms, cnt, ns = (ctypes.c_double * 32)(), (ctypes.c_uint64 * 32)(), ctypes.c_int32()
_cabi.check(lib.die_profile_end(32, ms, cnt, ctypes.byref(ns)))
if rank == 0:
    recs = [(t0[i], t1[i], sm[i], lib.die_profile_stage_name(st[i]).decode()) for i in range(n.value)]
    print("launches recorded", len(recs), "span us", max(r[1] for r in recs), "per step", max(r[1] for r in recs) / (N - 64))
    lo = recs[len(recs) // 2][0]
    for r in sorted(recs):
        if lo <= r[0] < lo + 500:
            print(f"{r[0]-lo:8.1f} {r[1]-lo:8.1f} {r[1]-r[0]:6.1f}  lane{r[2]} {r[3]}")
    for i in range(ns.value):
        if cnt[i]:
            print(lib.die_profile_stage_name(i).decode(), f"{1e3*ms[i]/cnt[i]:.1f} us x {cnt[i]/(N-64):.2f}/step")
dist.barrier()
dist.destroy_process_group()
