"""Where a multi-rank step spends its time: CUDA events at every phase boundary of ClipLoss forward + backward on real
ranks (bsc.loss.PHASE_MARKS), averaged over steps, per rank.
    python -m torch.distributed.run --nproc-per-node W --master-addr 127.0.0.1 tools/dist_phases.py [n_global] [steps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "bioscan-clip_b200"))
import torch
import torch.distributed as dist

import bench
import bsc
import bsc.loss as L

rank, world, lr = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
if world > 1:
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
nl = n // world
feats, labels = bench.synth_loss_rows(torch, n, rank * nl, nl, dev, torch.bfloat16)
crit = bsc.ClipLoss(rank=rank, world_size=world, precision="tc16") if world > 1 else \
    bsc.ContrastiveLoss(None, bench.SCALE, precision="tc16")
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
acc, order = {}, []
all_marks = []
for k in range(3 + steps):
    flush.zero_()
    L.PHASE_MARKS = marks = []
    leaves = [f.detach().requires_grad_(True) for f in feats]
    loss = crit(leaves[0], leaves[1], leaves[2], labels, bench.SCALE) if world > 1 else crit(*leaves, labels)
    loss.backward()
    L.PHASE_MARKS = None
    if k >= 3:
        all_marks.append(marks)  # no host synchronisation between steps: the host runs ahead like a training loop's
torch.cuda.synchronize()
for marks in all_marks:
    for (a, ea), (b, eb) in zip(marks[:-1], marks[1:]):
        if b not in acc:
            acc[b] = 0.0
            order.append(b)
        acc[b] += ea.elapsed_time(eb)
    acc["TOTAL"] = acc.get("TOTAL", 0.0) + marks[0][1].elapsed_time(marks[-1][1])
for r in range(world):
    if r == rank:
        print(f"rank {rank}: " + "  ".join(f"{b} {acc[b] / steps * 1e3:.0f}us" for b in order + ["TOTAL"]), flush=True)
    if world > 1:
        dist.barrier()
if world > 1:
    dist.destroy_process_group()
