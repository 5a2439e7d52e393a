#!/bin/bash
# Last GPU call of round 2: (1) whole GPU suite on the default build, (2) the two bank-scan candidates (build.py
# --variant v2 -DBSC_SCAN_V2=1, --variant v3 -DBSC_SCAN_V2=2): retrieval tests, then the 32-query scan probe of all three
# builds, (3) the whole suite on the best candidate that passed, (4) bench.py on that build, (5) one ncu capture of the
# 32-query scan of a 625k-row shard.  Every step under its own timeout; results under gpurun_out/f_*.
cd /root/repo; mkdir -p gpurun_out; O=gpurun_out; L=$PWD/bioscan-clip_b200/bsc
S=$O/f_summary.txt; : > $S
t0=$(date +%s)
stamp() { echo "[$(( $(date +%s) - t0 )) s] $*" | tee -a $S; }

timeout 420 python -m pytest tests -x -q -m gpu > $O/f_suite_a.log 2>&1; stamp "suite default rc=$? $(tail -1 $O/f_suite_a.log)"
timeout 200 python __graft_entry__.py --smoke > $O/f_smoke.log 2>&1; stamp "smoke rc=$? $(tail -1 $O/f_smoke.log)"
passed=""
for v in v2 v3; do
  BSC_LIB_PATH=$L/libbsc_hot_$v.so timeout 240 python -m pytest tests/test_gpu_retrieval.py tests/test_zz_no_alloc.py -x -q -m gpu > $O/f_retr_$v.log 2>&1
  rc=$?; stamp "retrieval tests $v rc=$rc $(tail -1 $O/f_retr_$v.log)"
  [ $rc = 0 ] && passed="$passed $v"
done
timeout 150 python tools/scan_probe.py 5000000 > $O/f_probe_a.txt 2>&1; stamp "probe default rc=$?"; cat $O/f_probe_a.txt | tee -a $S
for v in v2 v3; do
  BSC_LIB_PATH=$L/libbsc_hot_$v.so timeout 150 python tools/scan_probe.py 5000000 > $O/f_probe_$v.txt 2>&1; stamp "probe $v rc=$?"; cat $O/f_probe_$v.txt | tee -a $S
done
# the candidate that passed its tests and has the shortest search on the 625k-row shard, if shorter than the default's
best=$(python tools/ab/choose_scan.py $passed)
stamp "passed:$passed  chosen: '$best'"
if [ -n "$best" ]; then
  BSC_LIB_PATH=$L/libbsc_hot_$best.so timeout 420 python -m pytest tests -x -q -m gpu > $O/f_suite_$best.log 2>&1
  rc=$?; stamp "suite $best rc=$rc $(tail -1 $O/f_suite_$best.log)"
  [ $rc = 0 ] || best=""
fi
lib=$L/libbsc_hot.so; [ -n "$best" ] && lib=$L/libbsc_hot_$best.so
stamp "bench on $lib"
BSC_LIB_PATH=$lib timeout 600 python bench.py --gpus 1 > $O/f_bench_n1.json 2> $O/f_bench_n1.err; stamp "bench rc=$?"
tail -2 $O/f_bench_n1.err; cut -c1-1500 $O/f_bench_n1.json
python - <<'EOF' 2>&1 | tee -a gpurun_out/f_summary.txt
import json
try:
    d = json.load(open("gpurun_out/f_bench_n1.json"))
    print("ms_per_step", d["ms_per_step"], "e2e", d["e2e"]["value"], "check", d["check"]["ok"], d["check"]["grad_rel"])
    print("fp32x3", json.dumps(d.get("fp32x3"))[:1800])
    r = d.get("retrieval", {})
    print("retrieval", json.dumps({k: r.get(k) for k in ("stream32_bf16", "roofline", "check", "e2e")})[:1500])
except Exception as e:
    print("bench line unreadable:", e)
EOF
# ncu of the 32-query scan over a 625k-row shard (the 8-GPU share of the 5M bank) on the chosen build
BSC_LIB_PATH=$lib timeout 200 ncu --set full --import-source on --clock-control none -k 'regex:tc_sim_kernel' -s 2 -c 1 -f -o $O/f_scan625k python tools/prof_scan.py 625000 32 > $O/f_ncu_scan.log 2>&1
stamp "ncu rc=$?"
ncu -i $O/f_scan625k.ncu-rep --page raw --csv > $O/f_scan625k_raw.csv 2>/dev/null && python tools/ncu_summary.py $O/f_scan625k_raw.csv > $O/f_scan625k.txt 2>&1
stamp done
