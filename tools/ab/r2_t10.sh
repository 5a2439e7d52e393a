#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_loss.py tests/test_gpu_retrieval.py -x -q -m gpu 2>&1 | tail -3
python tools/time_step.py 32768 8
python tools/stripe_step.py 8 32768 8 scatter
python tools/scan_probe.py 2>&1 | tail -6
