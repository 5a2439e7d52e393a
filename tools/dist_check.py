"""torchrun check of ClipLoss on real GPUs against the ORACLE (fp64 closed form of the reference maths on the global
batch, evaluated on the host): every rank's loss and the gradients of its local rows, for every exchange path --
NVLink peer memory (fused gather + scatter backward), plain NCCL (gather + blocks backward), and the recomputing sweep.
Also: loss bits identical across ranks and across paths, and the autograd corner cases of the symmetric buffers
(two forwards before one backward, an evaluation forward in between).
    python -m torch.distributed.run --nproc-per-node W --master-addr 127.0.0.1 tools/dist_check.py [n_global] [d]"""
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
from bsc import _lib
from oracle import loss_oracle

TOL = 2e-3  # north_star: loss and gradients within 2e-3 relative in the 16-bit tensor-core mode

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
nl = n // world
feats_all, _ = bench.synth_loss_rows(torch, n, 0, n, dev, torch.bfloat16)
labels_all = torch.randint(0, n // 4, (n,), device=dev, generator=torch.Generator(device=dev).manual_seed(11))
sl = slice(rank * nl, (rank + 1) * nl)
# bf16-representable values in float32 leaves: the gradients come back un-rounded, the 2e-3 bound is the kernels' alone
local = [f[sl].float().contiguous() for f in feats_all]
local_labels = labels_all[sl].contiguous()

ref_in = [f.float().cpu() for f in feats_all]
if n <= 4096:
    ref_loss, ref_grads = loss_oracle.closed_form_grads(ref_in, labels_all.cpu(), bench.SCALE)
    ref_rows = torch.arange(n)
else:  # block-wise fp64 closed form on this rank's device, a strided sample of its rows
    ref_rows = torch.arange(rank * nl, (rank + 1) * nl, max(1, nl // 512))
    ref_loss, ref_grads = loss_oracle.blockwise_loss_and_grad_rows(ref_in, labels_all.cpu(), bench.SCALE,
                                                                   sample_rows=ref_rows, device=dev)
gmax = max(g.abs().max().item() for g in ref_grads)


def run(crit, feats, labels, *extra):
    leaves = [f.detach().clone().requires_grad_(True) for f in feats]
    loss = crit(leaves[0], leaves[1], leaves[2], labels, *extra)
    loss.backward()
    torch.cuda.synchronize()
    return loss.detach(), [l.grad for l in leaves]


def errors(loss, grads, ref_loss=ref_loss, ref_grads=ref_grads, gmax=gmax):
    dl = abs(float(loss) - float(ref_loss)) / abs(float(ref_loss))
    if n <= 4096:
        dg = max((g.double().cpu() - r[sl]).abs().max().item() for g, r in zip(grads, ref_grads))
    else:
        dg = max((g.double().cpu()[ref_rows - rank * nl] - r).abs().max().item() for g, r in zip(grads, ref_grads))
    return dl, dg / gmax


res = {}
for name, kw in (("nvlink", dict(fused_gather="on", backward="auto")), ("nccl", dict(fused_gather="off", backward="auto")),
                 ("nvlink-recompute", dict(fused_gather="on", backward="recompute"))):
    crit = bsc.ClipLoss(rank=rank, world_size=world, precision="tc16", **kw)
    for rep in range(3):  # several steps: the symmetric buffers alternate
        loss, grads = run(crit, local, local_labels, bench.SCALE)
    mode = _lib.last_bwd_mode()
    dist.barrier()
    t0 = time.perf_counter()
    for rep in range(10):
        loss, grads = run(crit, local, local_labels, bench.SCALE)
    dt = (time.perf_counter() - t0) / 10
    dl, dg = errors(loss, grads)
    res[name] = (float(loss), dl, dg, mode, grads)
    print(f"rank {rank} {name}: backward mode {mode} loss {float(loss):.7f} rel err {dl:.2e} "
          f"gradient max err / max|g| {dg:.2e}  {dt * 1e3:.3f} ms/step", flush=True)
    assert dl <= TOL and dg <= TOL, (name, dl, dg)
assert res["nvlink"][3] == _lib.BWD_SCATTER and res["nccl"][3] == _lib.BWD_BLOCKS and res["nvlink-recompute"][3] == 0
assert res["nvlink"][0] == res["nccl"][0] == res["nvlink-recompute"][0], "every path must give the same loss bits"
# the loss is the GLOBAL batch's on every rank: same bits everywhere
bits = torch.tensor([res["nvlink"][0]], dtype=torch.float64, device=dev)
lo, hi = bits.clone(), bits.clone()
dist.all_reduce(lo, op=dist.ReduceOp.MIN)
dist.all_reduce(hi, op=dist.ReduceOp.MAX)
assert float(lo) == float(hi), "ranks disagree on the loss"

# gather_with_grad=True: the reference's autograd-aware gather makes every rank's (identical) loss reach every row,
# i.e. world_size times the gradient (loss_func.py:81-82)
crit_g = bsc.ClipLoss(rank=rank, world_size=world, precision="tc16", gather_with_grad=True, fused_gather="on")
_, grads_g = run(crit_g, local, local_labels, bench.SCALE)
for g1, gw in zip(res["nvlink"][4], grads_g):
    assert torch.allclose(gw, g1 * world, rtol=1e-6, atol=0)

# autograd corner cases of the alternating symmetric buffers: two forwards (+ an evaluation forward) before the
# backwards, in both orders; the gradients must equal those of the plain fwd-bwd steps bit for bit
crit = bsc.ClipLoss(rank=rank, world_size=world, precision="tc16", fused_gather="on")
base = run(crit, local, local_labels, bench.SCALE)
other = [torch.roll(f, 1, dims=0) for f in local]
base_other = run(crit, other, local_labels, bench.SCALE)
la = [f.detach().clone().requires_grad_(True) for f in local]
lb = [f.detach().clone().requires_grad_(True) for f in other]
loss_a = crit(la[0], la[1], la[2], local_labels, bench.SCALE)
loss_b = crit(lb[0], lb[1], lb[2], local_labels, bench.SCALE)
with torch.no_grad():
    crit(la[0], la[1], la[2], local_labels, bench.SCALE)
    crit(lb[0], lb[1], lb[2], local_labels, bench.SCALE)
(loss_a + loss_b).backward()
torch.cuda.synchronize()
assert float(loss_a.detach()) == float(base[0]) and float(loss_b.detach()) == float(base_other[0])
for got, want in ((la, base[1]), (lb, base_other[1])):
    for leaf, g in zip(got, want):
        assert torch.equal(leaf.grad, g), "delayed backward read operands that a later forward had overwritten"

# float32-grade tensor-core mode (precision="fp32x3") on real ranks: full float32 inputs (not 16-bit representable), rows
# gathered with NCCL, scatter backward over the peer-mapped exchange buffers / blocks backward without any exchange;
# bound 1e-5 against the fp64 closed form
X3_TOL = 1e-5
feats32, _ = bench.synth_loss_rows(torch, n, 0, n, dev, torch.float32)
local32 = [f[sl].contiguous() for f in feats32]
if n <= 4096:
    ref32_loss, ref32_grads = loss_oracle.closed_form_grads([f.cpu() for f in feats32], labels_all.cpu(), bench.SCALE)
else:
    ref32_loss, ref32_grads = loss_oracle.blockwise_loss_and_grad_rows(feats32, labels_all, bench.SCALE,
                                                                       sample_rows=ref_rows, device=dev)
gmax32 = max(g.abs().max().item() for g in ref32_grads)
x3 = {}
for name, kw in (("fp32x3-scatter", dict(backward="auto")), ("fp32x3-blocks", dict(backward="blocks"))):
    crit = bsc.ClipLoss(rank=rank, world_size=world, precision="fp32x3", **kw)
    for rep in range(2):
        loss, grads = run(crit, local32, local_labels, bench.SCALE)
    dl, dg = errors(loss, grads, ref32_loss, ref32_grads, gmax32)
    x3[name] = (float(loss), _lib.last_bwd_mode())
    print(f"rank {rank} {name}: backward mode {_lib.last_bwd_mode()} loss {float(loss):.7f} rel err {dl:.2e} "
          f"gradient max err / max|g| {dg:.2e}", flush=True)
    assert dl <= X3_TOL and dg <= X3_TOL, (name, dl, dg)
assert x3["fp32x3-scatter"][1] == _lib.BWD_SCATTER and x3["fp32x3-blocks"][1] == _lib.BWD_BLOCKS
assert x3["fp32x3-scatter"][0] == x3["fp32x3-blocks"][0], "both strategies must give the same loss bits"
dist.barrier()
if rank == 0:
    print(f"dist_check ok: world {world} n {n}")
dist.destroy_process_group()
