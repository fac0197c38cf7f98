#!/bin/bash
# Round-2 profiling pass (run on the GPU box through gpurun; outputs under gpurun_out/).
# 1. launch list of the bench command, 2. --set full of the persistent step kernel inside that command,
# 3. --set full of the bf16 MLP GEMM and of the streaming TF32 layer (tensor-pipe evidence).
set -x
mkdir -p gpurun_out
BENCH="python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-kernels"
$BENCH > gpurun_out/prof_plain.log 2>&1 || exit 1
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r02_bench.csv $BENCH > gpurun_out/ncu_launches.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:pipe_step_kernel -s 30 -c 2 -f -o gpurun_out/prof_r02_pipe_step $BENCH > gpurun_out/ncu_full_step.log 2>&1
MLP='import numpy as np, torch, sys; sys.path[:0]=["distributed-inference-engine_b200"]; from die import ops
x=torch.randn(8192,4096,device="cuda").bfloat16(); w1=torch.randn(4096,4096,device="cuda").bfloat16(); w2=torch.randn(4096,4096,device="cuda").bfloat16(); b=torch.zeros(4096,device="cuda")
[ops.mlp_fwd_bf16(x,w1,b,w2,b) for _ in range(4)]
a=torch.randn(131072,128,device="cuda"); w=torch.randn(128,128,device="cuda"); bb=torch.zeros(128,device="cuda")
[ops.linear_fwd(a,w,bb,1) for _ in range(4)]; torch.cuda.synchronize()'
python -c "$MLP" > gpurun_out/prof_mlp_plain.log 2>&1 || exit 1
ncu --set full --clock-control none --import-source on -k regex:"gemm_bf16_kernel|linear_tf32_stream" -s 4 -c 3 -f -o gpurun_out/prof_r02_gemm python -c "$MLP" > gpurun_out/ncu_full_gemm.log 2>&1
ls -la gpurun_out/*.ncu-rep
