#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
W=${1:-2}
timeout 900 python -m pytest tests/test_gpu_loss.py -x -q -m gpu 2>&1 | tail -4
for n in 4096 32768; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29541 tools/dist_check.py $n 2>&1 | grep "^rank 0\|dist_check\|Error\|error\|Traceback\|bsc:" | head -20
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29544 tools/dist_phases.py 32768 10 2>&1 | grep "^rank\|Error\|error" | tee gpurun_out/r2e_phases_w$W.log
BSC_GATHER_OVERLAP=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29544 tools/dist_phases.py 32768 10 2>&1 | grep "^rank\|Error\|error" | tee gpurun_out/r2e_phases_nooverlap_w$W.log
