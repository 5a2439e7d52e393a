#!/bin/bash
for v in 1 0 1 0; do
  BSC_BENCH_NO_SAMPLER=$v timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 2 --no-retrieval --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('nosampler=$v', round(j['ms_per_step'],3), j['step_ms'])"
done
