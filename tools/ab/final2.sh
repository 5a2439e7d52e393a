set -x
cd /root/repo
W=${1:-2}
timeout 400 python -m pytest tests/test_gpu_multi.py -m gpu -x -q 2>&1 | tail -6 > gpurun_out/pytest_multi_w$W.log
tail -3 gpurun_out/pytest_multi_w$W.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $W > gpurun_out/bench_n$W.json 2> gpurun_out/bench_n$W.err
cat gpurun_out/bench_n$W.json
