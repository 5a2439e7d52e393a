#!/bin/bash
# bench.py on W GPUs of one box (W = 1: plain python), json + stderr under gpurun_out/
cd /root/repo
mkdir -p gpurun_out
W=${1:-1}; TAG=${2:-r2}; shift; shift
if [ "$W" = "1" ]; then
  timeout 900 python bench.py --gpus 1 "$@" > gpurun_out/${TAG}_bench_w1.json 2> gpurun_out/${TAG}_bench_w1.err
else
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus $W "$@" > gpurun_out/${TAG}_bench_w${W}.json 2> gpurun_out/${TAG}_bench_w${W}.err
fi
echo "bench rc=$?"; tail -3 gpurun_out/${TAG}_bench_w${W}.err; cat gpurun_out/${TAG}_bench_w${W}.json
