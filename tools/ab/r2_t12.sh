#!/bin/bash
cd /root/repo; mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_loss.py -x -q -m gpu -k "autograd_worker or config2 or error_behaviour" 2>&1 | tail -3
python tools/train_step.py 256 3 2>&1 | tail -2 | cut -c1-600
