"""One warm-up + N timed fwd+bwd steps of the bench workload (32768 x 768, 3 modalities); the command ncu wraps."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "bioscan-clip_b200"))
import torch  # noqa: E402

import bench  # noqa: E402
import bsc  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
torch.cuda.set_device(0)
feats, labels = bench.synth_loss_rows(torch, n, 0, n, torch.device("cuda", 0), torch.bfloat16)
crit = bsc.ContrastiveLoss(torch.nn.CrossEntropyLoss(), bench.SCALE, precision="tc16")
for _ in range(1 + steps):
    leaves = [f.detach().requires_grad_(True) for f in feats]
    loss = crit(leaves[0], leaves[1], leaves[2], labels)
    loss.backward()
torch.cuda.synchronize()
print("loss", float(loss.detach()))
