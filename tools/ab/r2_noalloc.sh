#!/bin/bash
# the no-allocation entry point: alone in a fresh process, then the whole GPU suite (it is collected last)
cd /root/repo; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu 2>&1 | tail -15 > gpurun_out/r2_noalloc_alone.log
cat gpurun_out/r2_noalloc_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -30 > gpurun_out/r2_suite.log
cat gpurun_out/r2_suite.log
