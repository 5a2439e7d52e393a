"""Bring-up helper run on the GPU box: each stage in its own process so one fault does not hide the rest.
Writes a readable log to gpurun_out/debug_*.log.  A checker like the tests (it calls the oracle), hence kept under tests/; not collected by pytest."""
import os
import subprocess
import sys
import textwrap

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT = os.path.join(ROOT, "gpurun_out")
os.makedirs(OUT, exist_ok=True)

PRE = textwrap.dedent(f"""
    import sys, os, time
    sys.path.insert(0, {ROOT!r}); sys.path.insert(0, os.path.join({ROOT!r}, 'bioscan-clip_b200'))
    import numpy as np, torch
    import bsc
    from bsc import _lib
    from bsc.ops import CudaOps, Problem
    from oracle import loss_oracle, retrieval_oracle as ro
    from tests.synth import loss_inputs, bank_inputs
    torch.cuda.set_device(0)
    ops = CudaOps()
    def rel(a, b):
        a = a.double().cpu(); b = b.double().cpu()
        return ((a - b).abs().max() / b.abs().max().clamp_min(1e-30)).item()
""")

STAGES = {}

STAGES["prep"] = """
x = torch.randn(300, 200, device='cuda') * 3
y = ops.normalize(x, torch.float32); print('l2norm f32 rel', rel(y, torch.nn.functional.normalize(x, dim=1)))
y = ops.normalize(x.bfloat16(), torch.float16); print('l2norm bf16->f16 rel', rel(y.float(), torch.nn.functional.normalize(x.bfloat16().float(), dim=1)))
lab = torch.randint(0, 50, (300,), device='cuda')
r, rep, nxt, rmax = ops.label_stats(lab)
t = (lab[None] == lab[:, None])
print('r ok', torch.equal(r, t.sum(1).float()), 'rmax', rmax.item(), t.sum(1).max().item())
print('rep ok', torch.equal(rep.long(), t.float().argmax(1)))
"""

STAGES["loss_fp32"] = """
for (n, d, pres, lab, nrm) in [(64, 32, (1,1,0), 'arange', True), (300, 200, (1,1,1), 'multi', False), (1024, 768, (1,1,1), 'arange', True)]:
    feats, labels = loss_inputs(n, d, tuple(bool(p) for p in pres), seed=n, normalised=nrm, labels=lab)
    ref_loss, ref_grads = loss_oracle.closed_form_grads(feats, labels)
    leaves = [None if f is None else f.cuda().requires_grad_(True) for f in feats]
    loss = bsc.ContrastiveLoss(None, 1/0.07, precision='fp32')(*leaves, labels.cuda())
    loss.backward(); torch.cuda.synchronize()
    print(n, d, 'loss', float(loss), float(ref_loss), 'rel', abs(float(loss)-float(ref_loss))/abs(float(ref_loss)))
    for l, g in zip(leaves, ref_grads):
        if l is not None: print('   grad rel', rel(l.grad, g))
"""

STAGES["tc_fwd"] = """
# forward statistics of the tcgen05 sweep vs a dense torch evaluation on the same fp16 operands
for (n, d) in [(128, 64), (256, 768), (1000, 768), (4096, 768)]:
    feats, labels = loss_inputs(n, d, seed=n, labels='multi', dtype=torch.bfloat16, device='cuda')
    prep = [ops.rows_to_f16(f) for f in feats]
    inv = [q[1] for q in prep]
    xh = [torch.nn.functional.normalize(f.double(), dim=1) for f in feats]
    r, rep, nxt, rmax = ops.label_stats(labels)
    p = Problem(xh=[q[0] for q in prep], n_global=n, n_local=n, row0=0, d=d, mode=_lib.MODE_TC16, scale=1/0.07, r=r, rep=rep, nxt=nxt, rmax=rmax, inv=inv)
    t0 = time.time(); stats = ops.infonce_fwd(p); torch.cuda.synchronize(); dt = time.time() - t0
    s = 1/0.07
    worst = 0
    for a in range(3):
        for b in range(3):
            if a == b: continue
            e = torch.exp(s * xh[a].double() @ xh[b].double().T - s).sum(1)
            pid = a*2 + (b-1 if b > a else b)
            got = stats[pid*n:(pid+1)*n].double()
            worst = max(worst, ((got - e).abs() / e).max().item())
    t = (labels[None] == labels[:, None]).double()
    pos = sum(((s * xh[a].double() @ xh[b].double().T - s) * t).sum() for a in range(3) for b in range(a + 1, 3))
    print(n, d, 'sumexp worst rel', worst, 'pos rel', abs(stats[6*n].item() - pos.item())/abs(pos.item()), 'time', dt)
"""

STAGES["tc_loss"] = """
for (n, d, pres, lab, sh) in [(256, 768, (1,1,0), 'arange', 0.5), (512, 768, (1,1,1), 'multi', 0.5), (1000, 768, (1,1,1), 'arange', 0.5), (384, 256, (1,0,1), 'multi', 0.0), (4096, 768, (1,1,1), 'arange', 0.5)]:
    feats, labels = loss_inputs(n, d, tuple(bool(p) for p in pres), seed=n+1, labels=lab, shared=sh, dtype=torch.bfloat16)
    ref_loss, ref_grads = loss_oracle.closed_form_grads([None if f is None else f.float() for f in feats], labels, dtype=torch.float32 if n > 2048 else torch.float64)
    for io in ('bf16', 'fp32'):
        leaves = [None if f is None else (f if io == 'bf16' else f.float()).cuda().requires_grad_(True) for f in feats]
        loss = bsc.ContrastiveLoss(None, 1/0.07, precision='tc16')(*leaves, labels.cuda())
        loss.backward(); torch.cuda.synchronize()
        print(n, d, io, 'loss', float(loss.detach()), float(ref_loss), 'rel', abs(float(loss.detach())-float(ref_loss))/abs(float(ref_loss)))
        for l, g in zip(leaves, ref_grads):
            if l is not None: print('   grad rel', rel(l.grad.float(), g))
"""

STAGES["pair_diag"] = """
for (n, pres) in [(128, (1,1,0)), (256, (1,1,0)), (512, (1,1,1))]:
    feats, labels = loss_inputs(n, 768, tuple(bool(p) for p in pres), seed=n+1, labels='arange', shared=0.5, dtype=torch.bfloat16)
    ref_loss, ref_grads = loss_oracle.closed_form_grads([None if f is None else f.float() for f in feats], labels)
    leaves = [None if f is None else f.float().cuda().requires_grad_(True) for f in feats]
    loss = bsc.ContrastiveLoss(None, 1/0.07, precision='tc16')(*leaves, labels.cuda())
    loss.backward(); torch.cuda.synchronize()
    for m, (l, g) in enumerate(zip(leaves, ref_grads)):
        if l is None: continue
        e = (l.grad.double().cpu() - g).abs(); sc = g.abs().max()
        print(n, 'mod', m, 'chunk0 err', (e[:, :384].max()/sc).item(), 'chunk1 err', (e[:, 384:].max()/sc).item(),
              'rowblk errs', [(e[i:i+128].max()/sc).item() for i in range(0, n, 128)])
        if n == 128 and m == 0:
            # undo the projection to look at the raw sweep output: dx = (ghat - xh (xh.ghat))/|x| ; compare the component orthogonal to xh
            o = l.grad.double().cpu(); xh = torch.nn.functional.normalize(feats[m].double(), dim=1)
            for blk in range(6):
                a = o[:, blk*128:(blk+1)*128]; b = g[:, blk*128:(blk+1)*128]
                print('      colblk', blk, 'corr', (a*b).sum().item() / (a.norm()*b.norm()).item(), 'norm ratio', (a.norm()/b.norm()).item())
            d = (o - g); print('      row err norms first 8 rows', [round(d[i, 384:].norm().item() / g[i, 384:].norm().item(), 3) for i in range(8)])
            print('      row err norms rows 64..72', [round(d[i, 384:].norm().item() / g[i, 384:].norm().item(), 3) for i in range(64, 72)])
        ratio = (l.grad.double().cpu() / g)
        print('    median ratio chunk0', ratio[:, :384].median().item(), 'chunk1', ratio[:, 384:].median().item())
"""

STAGES["retr_fp32"] = """
for (nb, nq, d, k) in [(333, 10, 40, 3), (5000, 257, 768, 5), (3, 4, 16, 5)]:
    bank, q, pi = bank_inputs(nb, nq, d, seed=nb)
    rs, ri = ro.search_ip(bank, q, k)
    ix = bsc.IndexFlatIP(d); ix.add(bank); s, i = ix.search(q, k)
    print(nb, nq, d, 'ids equal', np.array_equal(i, ri), 'score err', np.abs(s - rs).max())
"""

STAGES["retr_bf16"] = """
for (nb, nq, d, k) in [(777, 33, 128, 1), (4096, 128, 768, 5), (70000, 32, 768, 5)]:
    bank, q, pi = bank_inputs(nb, nq, d, seed=nb+1)
    rb = torch.from_numpy(bank).bfloat16().float().numpy(); rq = torch.from_numpy(q).bfloat16().float().numpy()
    rs, ri = ro.search_ip(rb, rq, k)
    ix = bsc.IndexFlatIP(d, precision='bf16'); ix.add(bank); s, i = ix.search(q, k)
    print(nb, nq, d, 'top1 equal', np.array_equal(i[:,0], ri[:,0]), 'all equal', np.array_equal(i, ri), 'score err', np.abs(s - rs).max())
"""

STAGES["perf"] = """
_lib.set_timing(True)
for n in (4096, 32768):
    feats, labels = loss_inputs(n, 768, seed=3, labels='arange', dtype=torch.bfloat16, device='cuda')
    crit = bsc.ContrastiveLoss(None, 1/0.07, precision='tc16')
    for it in range(3):
        leaves = [f.clone().requires_grad_(True) for f in feats]
        torch.cuda.synchronize(); t0 = time.time()
        loss = crit(*leaves, labels); loss.backward(); torch.cuda.synchronize(); dt = time.time() - t0
        print(n, 'step ms', dt*1e3, 'fwd kernel ms', _lib.last_kernel_ms(0), 'bwd kernel ms', _lib.last_kernel_ms(1), 'loss', float(loss))
    flops = 18 * n * n * 768
    print('  algorithmic TFLOP/s on fwd+bwd kernels:', flops / ((_lib.last_kernel_ms(0) + _lib.last_kernel_ms(1)) * 1e-3) / 1e12)
g = torch.Generator(device='cuda').manual_seed(3)
bank = torch.nn.functional.normalize(torch.randn(625000, 768, device='cuda', generator=g), dim=1)
ix = bsc.IndexFlatIP(768, precision='bf16'); ix.add(bank)
for nq in (32, 128, 4096):
    q = torch.nn.functional.normalize(torch.randn(nq, 768, device='cuda', generator=g), dim=1)
    for it in range(3):
        s, i = ix.search_device(q, 5); torch.cuda.synchronize()
    ms = _lib.last_kernel_ms(2)
    print('bank scan nq', nq, 'kernel ms', ms, 'GB/s', 625000*768*2/ms/1e6, 'TFLOP/s', 2*nq*625000*768/ms/1e9)
"""


def main():
    names = sys.argv[1:] or list(STAGES)
    for name in names:
        log = os.path.join(OUT, f"debug_{name}.log")
        with open(log, "w") as f:
            try:
                r = subprocess.run([sys.executable, "-c", PRE + STAGES[name]], stdout=f, stderr=subprocess.STDOUT,
                                   timeout=300)
                f.write(f"\n[exit {r.returncode}]\n")
            except subprocess.TimeoutExpired:
                f.write("\n[timeout]\n")
        print("=====", name)
        print(open(log).read()[-3000:])


if __name__ == "__main__":
    main()
