#!/bin/bash
# fp32x3 at the headline batch (gradient GEMMs cut into several launches) + the bench's fp32x3 block
cd /root/repo; mkdir -p gpurun_out; O=gpurun_out
timeout 300 python -m pytest tests/test_gpu_loss.py -q -s -m gpu -k "fp32x3" > $O/g_x3.log 2>&1; echo "x3 tests rc=$?"
grep -a "fp32x3 \|passed\|failed\|^FAILED\|^E  " $O/g_x3.log | cut -c1-300
timeout 300 python -m pytest tests/test_gpu_retrieval.py -q -m gpu > $O/g_retr.log 2>&1; echo "retrieval rc=$?"; tail -2 $O/g_retr.log
timeout 400 python bench.py --gpus 1 --no-retrieval > $O/g_bench.json 2> $O/g_bench.err; echo "bench rc=$?"; tail -2 $O/g_bench.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/g_bench.json"))
print("ms_per_step", d["ms_per_step"], "check", d["check"]["ok"])
print(json.dumps(d.get("fp32x3"), indent=1)[:3000])
PY
