"""Exact float32 search (screen + rescore) vs the bf16 scan and the exhaustive float32 kernel on a big bank.
    python tools/prof_exact.py [bank_rows] [queries]"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "bioscan-clip_b200"))
import torch
import bsc
nb = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 16384
torch.cuda.set_device(0); dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(21)
bank = torch.empty((nb, 768), dtype=torch.float32, device=dev)
for lo in range(0, nb, 262144):
    hi = min(nb, lo + 262144)
    bank[lo:hi] = torch.nn.functional.normalize(torch.randn(hi - lo, 768, generator=g, device=dev), dim=1)
pi = torch.randint(0, nb, (nq,), generator=g, device=dev)
q = torch.nn.functional.normalize(bank[pi] + 0.3 / 768 ** 0.5 * torch.randn(nq, 768, generator=g, device=dev), dim=1)
q[: nq // 100] = torch.nn.functional.normalize(torch.randn(nq // 100, 768, generator=g, device=dev), dim=1)  # 1 % random


def timed(index, queries, reps=2):
    index.search_device(queries, 5); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        s, i = index.search_device(queries, 5)
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps, s, i


ex = bsc.IndexFlatIP(768, precision="exact"); ex.add(bank)
t_ex, s_ex, i_ex = timed(ex, q)
print(f"exact (screen+rescore): {t_ex*1e3:.1f} ms for {nq} queries = {nq/t_ex:.0f} q/s, rechecked {ex.last_rechecked}, "
      f"top-1 planted hit {(i_ex[nq//100:, 0] == pi[nq//100:]).float().mean().item():.4f}")
b16 = bsc.IndexFlatIP(768, precision="bf16"); b16.add(bank)
t_b, _, i_b = timed(b16, q)
print(f"bf16 scan: {t_b*1e3:.1f} ms = {nq/t_b:.0f} q/s; top-5 sets equal to exact for {(i_b == i_ex).all(dim=1).float().mean().item():.4f} of the queries")
del b16
f32 = bsc.IndexFlatIP(768, precision="fp32"); f32._chunks, f32._bases, f32.ntotal = ex._chunks, ex._bases, ex.ntotal
sub = q[:256].contiguous()
t_f, _, i_f = timed(f32, sub, reps=1)
print(f"exhaustive fp32 kernel: {t_f*1e3:.1f} ms for 256 queries = {256/t_f:.0f} q/s; ids equal to exact mode: {bool((i_f == i_ex[:256]).all())}")
bad = (~(i_f == i_ex[:256]).all(dim=1)).nonzero().flatten()
print("mismatching queries:", bad.tolist()[:10], "of", int(bad.numel()))
s_ex256 = s_ex[:256]
_, s_f_full, _ = None, None, None
for b in bad.tolist()[:3]:
    print(b, "fp32 ", i_f[b].tolist())
    print(b, "exact", i_ex[b].tolist(), [round(v, 6) for v in s_ex[b].tolist()])
    sc = (bank[i_f[b]] @ q[b]).tolist(); print("   exact scores of the fp32 kernel's ids", [round(v, 6) for v in sc])
