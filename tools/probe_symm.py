import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
t = symm.empty(1024, dtype=torch.float32, device=torch.device("cuda", rank))
hdl = symm.rendezvous(t, dist.group.WORLD)
print(rank, "ptrs", [hex(p) for p in hdl.buffer_ptrs], "sig", len(hdl.signal_pad_ptrs), hdl.signal_pad_size, flush=True)
t.fill_(float(rank))
hdl.barrier(channel=0)
peer = hdl.get_buffer((rank + 1) % world, (1024,), torch.float32)
print(rank, "peer value", float(peer[0]), flush=True)
hdl.barrier(channel=0)
import time
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(200): hdl.barrier(channel=0)
torch.cuda.synchronize(); print(rank, "barrier us", (time.perf_counter() - t0) / 200 * 1e6, flush=True)
dist.destroy_process_group()
