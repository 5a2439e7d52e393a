"""torchrun check of ClipLoss on real GPUs: fused NVLink gather vs NCCL gather vs the single-process loss on the
same global batch (loss bit-identical across ranks, gradients of the local rows equal).
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tools/dist_check.py [n_global]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "bioscan-clip_b200"))
import torch
import torch.distributed as dist

import bench
import bsc

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
nl = n // world
feats_all, labels_all = bench.synth_loss_rows(torch, n, 0, n, dev, torch.bfloat16)
labels_all = torch.randint(0, n // 4, (n,), device=dev, generator=torch.Generator(device=dev).manual_seed(11))
sl = slice(rank * nl, (rank + 1) * nl)


def run(crit, feats, labels, *extra):
    leaves = [f.detach().clone().requires_grad_(True) for f in feats]
    loss = crit(leaves[0], leaves[1], leaves[2], labels, *extra)
    loss.backward()
    torch.cuda.synchronize()
    return loss.detach(), [l.grad for l in leaves]


ref_loss, ref_grads = run(bsc.ContrastiveLoss(None, bench.SCALE, precision="tc16"), feats_all, labels_all)
res = {}
for mode in ("off", "on"):
    crit = bsc.ClipLoss(rank=rank, world_size=world, precision="tc16", fused_gather=mode)
    local = [f[sl].contiguous() for f in feats_all]
    for rep in range(3):  # several steps: the two symmetric buffers alternate
        loss, grads = run(crit, local, labels_all[sl].contiguous(), bench.SCALE)
    dist.barrier()
    t0 = time.perf_counter()
    for rep in range(10):
        loss, grads = run(crit, local, labels_all[sl].contiguous(), bench.SCALE)
    dt = (time.perf_counter() - t0) / 10
    dl = abs(float(loss) - float(ref_loss))
    dg = max((g.float() - r[sl].float()).abs().max().item() for g, r in zip(grads, ref_grads))
    gm = max(r.float().abs().max().item() for r in ref_grads)
    res[mode] = (float(loss), dl, dg, gm, dt * 1e3)
    print(f"rank {rank} fused_gather={mode}: loss {float(loss):.7f} |dloss| {dl:.2e} max|dgrad| {dg:.2e} "
          f"(grad scale {gm:.2e}) {dt * 1e3:.3f} ms/step", flush=True)
assert res["on"][0] == res["off"][0], "fused and NCCL gathers must give the same loss bits"
# gradients come back in the input dtype (bf16: 2^-9 of their value), and the two runs tile the sweep differently
assert res["on"][1] <= 1e-6 * abs(float(ref_loss)) + 1e-7 and res["on"][2] <= 4e-3 * res["on"][3]
assert res["off"][2] <= 4e-3 * res["off"][3]
dist.barrier()
if rank == 0:
    print("dist_check ok")
dist.destroy_process_group()
