#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
W=${1:-8}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29544 tools/dist_phases.py 32768 10 2>&1 | grep "^rank" | tee gpurun_out/r2_phases_w$W.log
