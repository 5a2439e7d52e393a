#!/bin/bash
# round 2: real ranks. dist_check against the oracle at W = 2 (n = 4096 and 32768), then bench at 2 GPUs
cd /root/repo
mkdir -p gpurun_out
W=${1:-2}
for n in 4096 32768; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29541 tools/dist_check.py $n > gpurun_out/r2_dist_w${W}_$n.log 2>&1
  echo "rc=$?" >> gpurun_out/r2_dist_w${W}_$n.log
  grep -v "^W1\|^\*\*\*\|OMP_NUM" gpurun_out/r2_dist_w${W}_$n.log | tail -25
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus $W --no-retrieval > gpurun_out/r2_bench_w${W}.json 2> gpurun_out/r2_bench_w${W}.err
echo "bench rc=$?"; tail -5 gpurun_out/r2_bench_w${W}.err; cat gpurun_out/r2_bench_w${W}.json
