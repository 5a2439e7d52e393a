"""Multi-GPU parity on real devices (needs >= 2 GPUs; skipped otherwise): ClipLoss with the fused NVLink gather and
with NCCL collectives vs the single-process loss on the same global batch (tools/dist_check.py under torchrun)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs")
@pytest.mark.parametrize("n_global", [1024, 4096])
def test_cliploss_two_gpus_fused_and_nccl_gather(n_global):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(29600 + n_global % 97), os.path.join(ROOT, "tools", "dist_check.py"),
           str(n_global)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "dist_check ok" in r.stdout
