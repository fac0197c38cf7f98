#!/bin/bash
# A/B of one environment variable on one box: tools/ab_env.sh NGPU VAR v1 v2 ...
N=$1; VAR=$2; shift; shift
for v in "$@"; do
  env $VAR=$v timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 1500 --warmup 200 --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$VAR=$v', d['n_gpus'], round(d['value']/1e6,2), round(d['ms_per_step']*1000,1))
"
done
