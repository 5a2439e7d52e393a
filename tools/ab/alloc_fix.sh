set -x
cd /root/repo
W=4
python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $W --no-retrieval 2>gpurun_out/alloc_fix.err | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['ms_per_step'], d['step_ms'], d['e2e']['value'], d['e2e_step_ms'], d['clocks'])" > gpurun_out/alloc_fix.log
cat gpurun_out/alloc_fix.log
