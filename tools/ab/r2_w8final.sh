#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
W=${1:-8}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29541 tools/dist_check.py 4096 > gpurun_out/r2_dist_check_w${W}_4096.txt 2>&1
grep "^rank 0\|dist_check\|Error\|error\|Traceback\|bsc:" gpurun_out/r2_dist_check_w${W}_4096.txt | head
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29544 tools/dist_phases.py 32768 10 2>&1 | grep "^rank\|Error\|error\|bsc:" | tee gpurun_out/r2l_phases_w$W.log | head -3
bash tools/ab/r2_bench.sh $W r2l --no-retrieval | tail -1 | cut -c1-300
