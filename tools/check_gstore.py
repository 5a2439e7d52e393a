"""Backward modes on one GPU (BSC_BWD_GSTORE = 0 recomputing sweep, 1 stored spans, 2 weight matrix + GEMMs): gradients and times.
    python tools/check_gstore.py [n ...]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "bioscan-clip_b200"))
import torch
import bench, bsc
from bsc import _lib

sizes = [int(a) for a in sys.argv[1:]] or [640, 1000, 4096, 32768]
torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
_lib.set_timing(True)


def run(n, store, reps):
    os.environ["BSC_BWD_GSTORE"] = str(int(store))
    feats, labels = bench.synth_loss_rows(torch, n, 0, n, dev, torch.bfloat16)
    if n <= 4096:  # repeated labels: positives off the diagonal
        labels = labels // 3
    crit = bsc.ContrastiveLoss(torch.nn.CrossEntropyLoss(), bench.SCALE, precision="tc16")
    ta, tb, tt = [], [], []
    for k in range(reps):
        leaves = [x.detach().float().requires_grad_(True) for x in feats]
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        loss = crit(leaves[0], leaves[1], leaves[2], labels)
        loss.backward()
        e1.record()
        torch.cuda.synchronize()
        ta.append(_lib.last_kernel_ms(1)); tb.append(_lib.last_kernel_ms(3) if store else 0.0); tt.append(e0.elapsed_time(e1))
    del crit
    return float(loss.detach()), [l.grad.clone() for l in leaves], min(ta), min(tb), min(tt)


for n in sizes:
    reps = 4 if n >= 8192 else 2
    l0, g0, a0, _, t0 = run(n, 0, reps)
    print(f"n={n}: loss {l0:.6f}; recompute: sweep {a0:.3f} ms, step {t0:.3f} ms")
    for mode in (1, 2):
        l1, g1, a1, b1, t1 = run(n, mode, reps)
        worst = 0.0
        for m in range(3):
            scale = g0[m].abs().max().item()
            worst = max(worst, (g0[m] - g1[m]).abs().max().item() / scale)
        print(f"   mode {mode}: loss {l1:.6f}; max |dgrad| / max |grad| = {worst:.3e}; "
              f"first kernel {a1:.3f} + gemm {b1:.3f} ms, step {t1:.3f} ms")
    torch.cuda.empty_cache()
