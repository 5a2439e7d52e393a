#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
timeout 2400 python -m pytest tests -x -q -m gpu 2>&1 | tail -30 > gpurun_out/r2_suite.log
cat gpurun_out/r2_suite.log
