#!/bin/bash
# usage: scale_n.sh N  -> runs own + reference arms at N GPUs (on the box)
N=$1
mkdir -p gpurun_out
for arm in own ref; do
  extra=""; [ $arm = ref ] && extra="--impl reference"
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 5 $extra > gpurun_out/scale_n${N}_$arm.json 2> gpurun_out/scale_n${N}_$arm.err
  echo "N=$N $arm rc=$?"; cut -c90-230 gpurun_out/scale_n${N}_$arm.json
done
