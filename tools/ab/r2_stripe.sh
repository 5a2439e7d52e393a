#!/bin/bash
# one rank's share of an 8-rank / 2-rank step replayed on one GPU: kernel times, then the ncu launch list
cd /root/repo
mkdir -p gpurun_out
for w in 8 2; do for m in scatter blocks; do python tools/stripe_step.py $w 32768 8 $m; done; done 2>&1 | tee gpurun_out/r2_stripe_time.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_stripe8.csv python tools/stripe_step.py 8 32768 2 scatter > gpurun_out/r2_ncu_stripe8.log 2>&1
tail -2 gpurun_out/r2_ncu_stripe8.log
