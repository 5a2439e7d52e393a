#!/bin/bash
# single-GPU step: launch list + full capture of the three sweeps
cd /root/repo; mkdir -p gpurun_out
python tools/time_step.py 32768 5
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file gpurun_out/r2_launches_step.csv python tools/prof_step.py 32768 2 > gpurun_out/r2_ncu_step.log 2>&1
tail -1 gpurun_out/r2_ncu_step.log
timeout 600 ncu --set full --import-source on --clock-control none -k 'regex:tc_sim_kernel|tc_bwd_rev' -s 3 -c 3 -f -o gpurun_out/r2_step_full python tools/prof_step.py 32768 1 > gpurun_out/r2_ncu_full.log 2>&1
tail -1 gpurun_out/r2_ncu_full.log
ncu -i gpurun_out/r2_step_full.ncu-rep --page raw --csv > gpurun_out/r2_step_full_raw.csv
ls -la gpurun_out/r2_step_full*
