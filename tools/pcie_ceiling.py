"""Diagnostic: what the box's PCIe link gives for the e2e step's traffic pattern (pinned host buffers, copy-in and
copy-out running at the same time on two streams), to put die_pipeline_run_host_many's e2e number in proportion."""
import time
import torch

MB = 1 << 20
n_in, n_out = 276 * MB, 269 * MB
hin = torch.empty(n_in, dtype=torch.uint8).pin_memory()
hout = torch.empty(n_out, dtype=torch.uint8).pin_memory()
din = torch.empty(n_in, dtype=torch.uint8, device="cuda")
dout = torch.empty(n_out, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()


def run(chunks, both=True, reps=10):
    ci, co = n_in // chunks, n_out // chunks
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(reps):
        for c in range(chunks):
            with torch.cuda.stream(s1):
                din[c * ci:(c + 1) * ci].copy_(hin[c * ci:(c + 1) * ci], non_blocking=True)
            if both:
                with torch.cuda.stream(s2):
                    hout[c * co:(c + 1) * co].copy_(dout[c * co:(c + 1) * co], non_blocking=True)
    torch.cuda.synchronize()
    return (time.perf_counter() - t) / reps


for chunks in (1, 8, 32):
    a = run(chunks, both=False)
    b = run(chunks, both=True)
    print(f"chunks {chunks:3d}: H2D alone {n_in / a / 1e9:6.1f} GB/s ({a * 1e3:.2f} ms) | H2D + D2H together {b * 1e3:.2f} ms per step "
          f"= {n_in / b / 1e9:.1f} + {n_out / b / 1e9:.1f} GB/s")
