"""Diagnostic: device time of the bf16 MLP forward (two tcgen05 GEMMs) at the C5 shape."""
import sys
import torch
sys.path[:0] = ["distributed-inference-engine_b200", "."]
from die import ops
B, D, H = 8192, 4096, 4096
x = torch.randn(B, D, device="cuda").bfloat16()
w1 = (torch.randn(H, D, device="cuda") / 64).bfloat16()
w2 = (torch.randn(D, H, device="cuda") / 64).bfloat16()
b1, b2 = torch.zeros(H, device="cuda"), torch.zeros(D, device="cuda")
for _ in range(3):
    ops.mlp_fwd_bf16(x, w1, b1, w2, b2)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
n = 20
e0.record()
for _ in range(n):
    ops.mlp_fwd_bf16(x, w1, b1, w2, b2)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / n
print("mlp forward %.1f us  %.0f TFLOP/s" % (ms * 1e3, 2 * 2 * B * D * H / ms / 1e9))
