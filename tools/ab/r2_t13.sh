#!/bin/bash
# the fp32x3 tensor-core mode: its tests alone, with the measured errors printed
cd /root/repo; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_loss.py -q -s -m gpu -k "fp32x3" > gpurun_out/r2_x3_full.log 2>&1
grep -a "fp32x3 \|passed\|failed\|^FAILED\|^E  " gpurun_out/r2_x3_full.log | cut -c1-330 > gpurun_out/r2_x3_tests.log
cat gpurun_out/r2_x3_tests.log
