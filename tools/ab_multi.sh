#!/bin/bash
# A/B of library builds on one box: tools/ab_multi.sh NGPU lib1 lib2 ... (each run twice, interleaved)
N=$1; shift
for rep in 1 2; do
for lib in "$@"; do
  DIE_LIB_PATH=$PWD/$lib timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 1500 --warmup 200 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$lib', d['n_gpus'], round(d['value']/1e6,2), round(d['ms_per_step']*1000,1), d['gpu_launches'])
"
done
done
