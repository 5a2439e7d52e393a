#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
W=${1:-8}; TAG=${2:-r2f}
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29544 tools/dist_phases.py 32768 10 2>&1 | grep "^rank [01]:\|Error\|error" | tee gpurun_out/${TAG}_phases_w$W.log
BSC_GATHER_OVERLAP=0 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29544 tools/dist_phases.py 32768 10 2>&1 | grep "^rank [01]:\|Error\|error" | tee gpurun_out/${TAG}_phases_nooverlap_w$W.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus $W --no-retrieval > gpurun_out/${TAG}_bench_w${W}.json 2> gpurun_out/${TAG}_bench_w${W}.err
echo "bench rc=$?"; tail -3 gpurun_out/${TAG}_bench_w${W}.err; python - <<PY
import json
z=json.load(open("gpurun_out/${TAG}_bench_w${W}.json"))
r=z["roofline"]
print("ms_per_step", z["ms_per_step"], "value", z["value"], "e2e", z["e2e"]["value"], "fwd", r["fwd_kernel_ms"], "bwd", r["backward_kernels_ms"], "step_frac", r["step_frac"], "steps", z["step_ms"], "check", z["check"]["ok"], z["check"]["grad_rel"], z["gather"])
PY
