#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
W=${1:-8}; B=${2:-512}
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29547 tools/train_step.py $B 5 2> gpurun_out/r2_train_step_n$W.err | grep "^{" | tee gpurun_out/r2_train_step_n$W.json
tail -5 gpurun_out/r2_train_step_n$W.err | cut -c1-300
