#!/bin/bash
# last call of the round: GPU suite, smoke and the bench line of the committed state
cd /root/repo; mkdir -p gpurun_out; O=gpurun_out
timeout 400 python -m pytest tests -x -q -m gpu > $O/z_suite.log 2>&1; echo "suite rc=$? $(tail -2 $O/z_suite.log | head -1)"
timeout 150 python __graft_entry__.py --smoke 2>&1 | tail -1
timeout 400 python bench.py --gpus 1 > $O/z_bench_n1.json 2> $O/z_bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d = json.load(open("gpurun_out/z_bench_n1.json"))
print("ms_per_step", d["ms_per_step"], "value", d["value"], "e2e", d["e2e"]["value"], "check", d["check"]["ok"], "clocks", d["clocks"])
print("x3", {k: (v.get("ms_per_step"), v.get("check", {}).get("grad_rel")) for k, v in d["fp32x3"].items() if isinstance(v, dict)})
r = d["retrieval"]
print("retr", r["bulk_exact_fp32"]["queries_per_s"], r["stream32_bf16"]["scan_kernel_ms"], r["roofline"]["frac"], r["check"]["ids_equal"], r["e2e"]["value"])
PY
