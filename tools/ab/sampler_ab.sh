set -x
cd /root/repo
W=4
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $W --steps 20 --no-retrieval 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('$1', d['ms_per_step'], d['step_ms'], d['clocks'])"; }
BSC_BENCH_NO_SAMPLER=1 run nosampler >> gpurun_out/sampler_ab.log
run sampler >> gpurun_out/sampler_ab.log
BSC_BENCH_NO_SAMPLER=1 run nosampler >> gpurun_out/sampler_ab.log
run sampler >> gpurun_out/sampler_ab.log
cat gpurun_out/sampler_ab.log
