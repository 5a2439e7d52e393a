"""Synthetic full train step (BASELINE config 5) on one GPU or under torchrun: LoRA ViT-B/16 + BarcodeBERT + BERT-small
with the fused loss; prints the step time and the share spent in the loss (CUDA events).
    python tools/train_step.py [batch_per_gpu] [steps]
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/train_step.py 512"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "bioscan-clip_b200"))
import torch
import torch.distributed as dist

import bsc
from bsc import harness

world, rank, lr = int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("RANK", 0)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
if world > 1:
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
bpg = int(sys.argv[1]) if len(sys.argv) > 1 else 256
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 5
torch.manual_seed(0)  # same initial weights on every rank (broadcast_model, train_cl.py:29-31)
model = harness.build_synthetic_model().to(dev)
opt = torch.optim.AdamW([p for p in model.parameters() if p.requires_grad], lr=1e-4)
if world > 1:
    crit, extra = bsc.ClipLoss(rank=rank, world_size=world, criterion=torch.nn.CrossEntropyLoss(), precision="tc16"), (1 / 0.07,)
else:
    crit, extra = bsc.ContrastiveLoss(torch.nn.CrossEntropyLoss(), 1 / 0.07, precision="tc16"), ()
batch = harness.synthetic_batch(bpg, dev, seed=rank)
batch = batch[:3] + (batch[3] + rank * bpg,)

t_loss = []


class Timed(torch.nn.Module):
    def forward(self, *a):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = crit(*a)
        e1.record()
        t_loss.append((e0, e1))
        return out


for _ in range(2):
    loss = harness.train_step(model, Timed(), opt, batch, group=dist.group.WORLD if world > 1 else None, loss_args=extra)
torch.cuda.synchronize()
t_loss.clear()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(steps):
    loss = harness.train_step(model, Timed(), opt, batch, group=dist.group.WORLD if world > 1 else None, loss_args=extra)
b.record()
b.synchronize()
ms = a.elapsed_time(b) / steps
fwd_loss_ms = sum(x.elapsed_time(y) for x, y in t_loss) / steps
# the loss alone (forward + backward) on embeddings of the same shape and dtype as the encoders hand over
with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
    outs = model(*batch[:3])
c, d2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for k in range(2 + steps):
    if k == 2:
        c.record()
    leaves = [o.detach().requires_grad_(True) for o in outs]
    crit(*leaves, batch[3], *extra).backward()
d2.record()
d2.synchronize()
loss_ms = c.elapsed_time(d2) / steps
if rank == 0:
    print(json.dumps({"workload": "LoRA ViT-B/16 + BarcodeBERT + BERT-small train step, synthetic, bf16 autocast",
                      "n_gpus": world, "batch_per_gpu": bpg, "global_batch": bpg * world, "ms_per_step": ms,
                      "samples_per_s": bpg * world / (ms * 1e-3), "loss_forward_ms_inside_the_step": fwd_loss_ms,
                      "loss_fwd_bwd_ms_alone": loss_ms, "loss_share_of_step": loss_ms / ms,
                      "embedding_dtype": str(outs[0].dtype), "loss": float(loss)}))
if world > 1:
    dist.destroy_process_group()
