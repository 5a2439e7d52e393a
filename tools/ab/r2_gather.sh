#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
W=${1:-8}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29545 tools/gather_probe.py 32768 2>&1 | grep "gather of\|Error\|error\|Traceback" | tee gpurun_out/r2_gather_w$W.log
