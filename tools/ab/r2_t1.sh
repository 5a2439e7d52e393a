#!/bin/bash
# round 2, first GPU pass: the refactored backward (plan / partials / finish, scatter strategy) on one GPU
cd /root/repo
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_loss.py -x -q -m gpu 2>&1 | tail -40 > gpurun_out/r2_t1_pytest.log
cat gpurun_out/r2_t1_pytest.log
