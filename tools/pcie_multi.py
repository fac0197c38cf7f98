"""Diagnostic: host <-> device copy bandwidth when every rank copies at once (what bounds e2e at N > 1), with and
without pinning each rank's CPU affinity to its GPU's NUMA node before the pinned buffers are allocated."""
import os
import sys
import time

import torch
import torch.distributed as dist

rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", "0"))
if os.environ.get("BIND") == "1":
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    import bench
    print(rank, "bound to", bench.bind_to_gpu_numa_node(local), flush=True)
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
MB = 1 << 20
n = 256 * MB
hin = torch.empty(n, dtype=torch.uint8).pin_memory()
hout = torch.empty(n, dtype=torch.uint8).pin_memory()
hin.fill_(1)
hout.fill_(1)
din = torch.empty(n, dtype=torch.uint8, device="cuda")
dout = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for both in (False, True):
    torch.cuda.synchronize()
    dist.barrier()
    t = time.perf_counter()
    for _ in range(8):
        with torch.cuda.stream(s1):
            din.copy_(hin, non_blocking=True)
        if both:
            with torch.cuda.stream(s2):
                hout.copy_(dout, non_blocking=True)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t) / 8
    print(f"rank {rank}/{world} {'H2D+D2H' if both else 'H2D only'}: {n / dt / 1e9:.1f} GB/s per direction", flush=True)
    dist.barrier()
dist.destroy_process_group()
