set -x
cd /root/repo
python tools/stripe_step.py 8 32768 8 > gpurun_out/stripe8_time.log 2>&1
python tools/stripe_step.py 2 32768 8 >> gpurun_out/stripe8_time.log 2>&1
cat gpurun_out/stripe8_time.log
timeout 200 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/launches_stripe8.csv python tools/stripe_step.py 8 32768 2 > gpurun_out/ncu_stripe8.log 2>&1
tail -2 gpurun_out/ncu_stripe8.log
timeout 300 ncu --set full --clock-control none -k regex:"tc_sim_kernel|tc_bwd_rev" -s 14 -c 3 -f -o gpurun_out/stripe8_full python tools/stripe_step.py 8 32768 1 > gpurun_out/ncu_stripe8_full.log 2>&1
tail -2 gpurun_out/ncu_stripe8_full.log
ncu -i gpurun_out/stripe8_full.ncu-rep --page raw --csv > gpurun_out/stripe8_full_raw.csv 2>/dev/null
ls -la gpurun_out/
