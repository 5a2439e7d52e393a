#!/bin/bash
# two real ranks: every exchange path incl. fp32x3 against the oracle, then the bench line (new few-query scan per shard)
cd /root/repo; mkdir -p gpurun_out; O=gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 tools/dist_check.py 4096 > $O/w2_dist_check.log 2>&1; echo "dist_check rc=$?"
grep -a "^rank\|dist_check ok\|Error\|assert" $O/w2_dist_check.log | cut -c1-220 | tail -24
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 2 > $O/w2_bench.json 2> $O/w2_bench.err; echo "bench rc=$?"; tail -2 $O/w2_bench.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/w2_bench.json"))
print("ms_per_step", d["ms_per_step"], "value", d["value"], "e2e", d["e2e"]["value"], "check", d["check"]["ok"], d["check"]["grad_rel"], d["gather"])
r = d.get("retrieval", {})
print("retrieval", json.dumps({k: r.get(k) for k in ("bulk_exact_fp32", "stream32_bf16", "roofline", "check", "e2e")})[:2200])
PY
