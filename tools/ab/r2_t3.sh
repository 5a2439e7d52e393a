#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_retrieval.py -x -q -m gpu 2>&1 | tail -15
python tools/scan_probe.py 2>&1 | tee gpurun_out/r2_scan_probe.log
