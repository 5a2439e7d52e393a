#!/bin/bash
for m in 3 0; do
  echo "== BSC_BWD_GSTORE=$m dist_check 4096"; BSC_BWD_GSTORE=$m timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29521 tools/dist_check.py 4096 2>&1 | grep "rank 0"
  echo "== BSC_BWD_GSTORE=$m bench"; BSC_BWD_GSTORE=$m timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 2 --no-retrieval --no-cpu-baseline 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); r=j['roofline']
print(round(j['ms_per_step'],3), round(j['value']), round(j['e2e']['value']), r['backward_mode'], r['backward_kernels_ms'], r['fwd_kernel_ms'])"
done
