#!/bin/bash
# final build, one GPU: launch list of the step (ncu durations), full capture of the three sweeps and of the 32-query
# scan over the 5M-row bank; each program first run without ncu
cd /root/repo; mkdir -p gpurun_out; O=gpurun_out
timeout 120 python tools/time_step.py 32768 5 | tail -1
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/p_launches_step.csv python tools/prof_step.py 32768 2 > $O/p_ncu_step.log 2>&1; echo "launch list rc=$?"
timeout 400 ncu --set full --import-source on --clock-control none -k 'regex:tc_sim_kernel|tc_bwd_rev' -s 3 -c 3 -f -o $O/p_step_full python tools/prof_step.py 32768 1 > $O/p_ncu_full.log 2>&1; echo "full capture rc=$?"
ncu -i $O/p_step_full.ncu-rep --page raw --csv > $O/p_step_full_raw.csv && python tools/ncu_summary.py $O/p_step_full_raw.csv > $O/p_step_kernels.txt
timeout 120 python tools/prof_scan.py 5000000 32 | tail -2
timeout 300 ncu --set full --clock-control none -k 'regex:tc_sim_kernel' -s 2 -c 1 -f -o $O/p_scan5m python tools/prof_scan.py 5000000 32 > $O/p_ncu_scan.log 2>&1; echo "scan capture rc=$?"
ncu -i $O/p_scan5m.ncu-rep --page raw --csv > $O/p_scan5m_raw.csv && python tools/ncu_summary.py $O/p_scan5m_raw.csv > $O/p_scan5m.txt
rm -f $O/p_step_full.ncu-rep $O/p_scan5m.ncu-rep
grep -a "duration\|tensor_cycles\|dram__bytes" $O/p_step_kernels.txt $O/p_scan5m.txt | cut -c1-160
