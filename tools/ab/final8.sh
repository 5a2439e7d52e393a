set -x
cd /root/repo
W=${1:-8}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $W > gpurun_out/bench_n$W.json 2> gpurun_out/bench_n$W.err
cat gpurun_out/bench_n$W.json
