#!/bin/bash
# two ranks: the bench's retrieval block through bsc.merge_across_ranks (one packed all_gather + merge)
cd /root/repo; mkdir -p gpurun_out; O=gpurun_out
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 2 --steps 5 > $O/w2m_bench.json 2> $O/w2m_bench.err; echo "bench rc=$?"; tail -2 $O/w2m_bench.err
python - <<'PY'
import json
d = json.load(open("gpurun_out/w2m_bench.json"))
print("ms_per_step", d["ms_per_step"], "check", d["check"]["ok"])
r = d.get("retrieval", {})
print("retrieval", json.dumps({k: r.get(k) for k in ("error", "bulk_exact_fp32", "stream32_bf16", "check", "e2e")})[:1800])
PY
