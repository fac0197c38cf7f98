"""Diagnostic: per-phase time of the peer-store dispatch on N GPUs (torchrun)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "distributed-inference-engine_b200")]
import numpy as np, torch, torch.distributed as dist
import bench
from die import ops, _cabi
from die.exchange import PeerDispatcher
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
B, D = bench.B, bench.D
w, b = bench.model_weights()
disp = PeerDispatcher(world, rank, 1 << 18, B, D, w, b, 300.0, model="linear_tf32")
rng = np.random.default_rng(rank)
off_d = torch.from_numpy(np.arange(B + 1, dtype=np.int64) * 15).cuda()
N = 100
_, keys, _, x = bench.make_batches(rng, N)
kd, xd = torch.from_numpy(keys).cuda(), torch.from_numpy(x).cuda()
kbs = [ops.KeyBatch(kd[i], off_d) for i in range(N)]; xs = [xd[i] for i in range(N)]
out = torch.empty((B, D), device="cuda"); hit = torch.empty(B, dtype=torch.uint8, device="cuda")
for i in range(20): disp.step(kbs[i], xs[i], 1.0 + i, out, hit)
torch.cuda.synchronize(); dist.barrier()
t0 = time.perf_counter()
for i in range(20, N): disp.step(kbs[i], xs[i], 1.0 + i, out, hit)
t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
print(rank, f"enqueue {1e6*(t1-t0)/(N-20):.1f} us/step total {1e6*(t2-t0)/(N-20):.1f} us/step", flush=True)
# phase split with syncs
lib = _cabi.load()
import ctypes
_cabi.check(lib.die_profile_begin())
for i in range(20): disp.step(kbs[i], xs[i], 200.0 + i, out, hit)
ms, cnt, n = (ctypes.c_double * 32)(), (ctypes.c_uint64 * 32)(), ctypes.c_int32()
_cabi.check(lib.die_profile_end(32, ms, cnt, ctypes.byref(n)))
if rank == 0:
    for i in range(n.value):
        if cnt[i]: print(lib.die_profile_stage_name(i).decode(), f"{1e3*ms[i]/cnt[i]:.1f} us x {cnt[i]/20:.1f}/step")
    hdl = disp.hdl
torch.cuda.synchronize(); dist.barrier()
t0 = time.perf_counter()
for _ in range(100): disp.hdl.barrier(channel=0)
torch.cuda.synchronize()
if rank == 0: print("barrier", (time.perf_counter() - t0) / 100 * 1e6, "us")
dist.destroy_process_group()
