#!/bin/bash
# whole GPU suite + bench of the current default build
cd /root/repo; mkdir -p gpurun_out; O=gpurun_out
timeout 500 python -m pytest tests -x -q -m gpu > $O/c_suite.log 2>&1; echo "suite rc=$? $(tail -2 $O/c_suite.log | head -1)"
grep -a "^FAILED\|^E  " $O/c_suite.log | cut -c1-250 | head
timeout 200 python -m pytest tests/test_gpu_loss.py -q -s -m gpu -k "fp32x3" 2>&1 | grep -a "fp32x3 \|passed\|failed" | cut -c1-220
timeout 500 python bench.py --gpus 1 > $O/c_bench_n1.json 2> $O/c_bench_n1.err; echo "bench rc=$?"; tail -2 $O/c_bench_n1.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/c_bench_n1.json"))
print("ms_per_step", d["ms_per_step"], "e2e", d["e2e"]["value"], "check", d["check"])
print("fp32x3", json.dumps(d.get("fp32x3"))[:2500])
r = d.get("retrieval", {})
print("retrieval", json.dumps({k: r.get(k) for k in ("bulk_exact_fp32", "stream32_bf16", "roofline", "check", "e2e")})[:2500])
PY
