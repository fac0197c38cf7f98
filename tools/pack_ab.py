import os, sys, numpy as np, torch
sys.path[:0]=["distributed-inference-engine_b200","."]
from die import ops
rng=np.random.default_rng(0)
n_req=64*256
data=[]
for r in range(3):
    lens=rng.integers(16,2049,n_req); off=np.zeros(n_req+1,dtype=np.int64); np.cumsum(lens,out=off[1:])
    vals=torch.empty(int(off[-1])+8,dtype=torch.float32,device="cuda").normal_()
    data.append((vals,torch.from_numpy(off).cuda(),int(off[-1])))
def timed(fn,n=30):
    for i in range(3): fn(i)
    torch.cuda.synchronize()
    e0,e1=torch.cuda.Event(enable_timing=True),torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n): fn(i)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1)*1e-3/n
outs=[torch.empty((n_req,2048),device="cuda") for _ in range(3)]
lib=ops.require_cuda() if hasattr(ops,"require_cuda") else None
packed=[ops.pack_ragged(d[0],d[1],2048)[0] for d in data]
tot=sum(d[2] for d in data)/3
s=timed(lambda i: ops.pack_ragged(data[i%3][0],data[i%3][1],2048))
print("pack   %.1f us  %.0f GB/s  frac %.3f"%(s*1e6,(4*tot+8*n_req+4*2048*n_req)/s/1e9,(4*tot+8*n_req+4*2048*n_req)/s/1e9/6554))
flat=[torch.empty_like(d[0]) for d in data]
s=timed(lambda i: ops.unpack_ragged(packed[i%3],data[i%3][1],flat[i%3]))
print("unpack %.1f us  %.0f GB/s  frac %.3f"%(s*1e6,(8*tot+8*n_req)/s/1e9,(8*tot+8*n_req)/s/1e9/6554))
for i in range(3): assert torch.equal(flat[i][:data[i][2]], data[i][0][:data[i][2]])
