"""Kernel times of the bench workload (CUDA events inside the library) and the time per step of 10 steps queued back to
back (the bench's timed loop without its extras): python tools/time_step.py [n] [steps]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "bioscan-clip_b200"))
import torch
import bench, bsc
from bsc import _lib
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
torch.cuda.set_device(0)
feats, labels = bench.synth_loss_rows(torch, n, 0, n, torch.device("cuda", 0), torch.bfloat16)
crit = bsc.ContrastiveLoss(torch.nn.CrossEntropyLoss(), bench.SCALE, precision="tc16")
_lib.set_timing(True)
f, b, g = [], [], []


def step():
    leaves = [x.detach().requires_grad_(True) for x in feats]
    loss = crit(leaves[0], leaves[1], leaves[2], labels)
    loss.backward()
    return loss, leaves


for k in range(2 + steps):
    loss, leaves = step()
    torch.cuda.synchronize()
    if k >= 2:
        f.append(_lib.last_kernel_ms(0)); b.append(_lib.last_kernel_ms(1)); g.append(_lib.last_kernel_ms(3))
_lib.set_timing(False)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(10)]
for a, e in ev:
    flush.zero_()
    a.record()
    keep = step()
    e.record()
torch.cuda.synchronize()
ms = [a.elapsed_time(e) for a, e in ev]
print("loss %.8f fwd_ms %.3f weights_ms %.3f gemm_ms %.3f | queued steps: mean %.3f ms  (%s)" % (
    float(loss.detach()), sum(f) / len(f), sum(b) / len(b), sum(g) / len(g),
    sum(ms) / len(ms), " ".join("%.2f" % v for v in ms)))
