"""Which of {exact mode, exhaustive fp32 kernel} agrees with a float64 torch evaluation at 5M rows?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "bioscan-clip_b200"))
import torch
import bsc
nb, nq = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000, 256
torch.cuda.set_device(0); dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(21)
bank = torch.empty((nb, 768), dtype=torch.float32, device=dev)
for lo in range(0, nb, 262144):
    hi = min(nb, lo + 262144)
    bank[lo:hi] = torch.nn.functional.normalize(torch.randn(hi - lo, 768, generator=g, device=dev), dim=1)
q = torch.nn.functional.normalize(torch.randn(nq, 768, generator=g, device=dev), dim=1)
best_s = torch.full((nq, 8), -1e30, dtype=torch.float64, device=dev); best_i = torch.full((nq, 8), -1, dtype=torch.int64, device=dev)
for lo in range(0, nb, 262144):
    hi = min(nb, lo + 262144)
    s = q.double() @ bank[lo:hi].double().T
    ts, ti = torch.topk(s, 8, dim=1)
    cs, ci = torch.cat([best_s, ts], 1), torch.cat([best_i, ti + lo], 1)
    ms, mi = torch.topk(cs, 8, dim=1)
    best_s, best_i = ms, torch.gather(ci, 1, mi)
print("min gap between rank 5 and 6 (float64):", (best_s[:, 4] - best_s[:, 5]).min().item())
ex = bsc.IndexFlatIP(768, precision="exact"); ex.add(bank)
s_ex, i_ex = ex.search_device(q, 5)
f32 = bsc.IndexFlatIP(768, precision="fp32"); f32._chunks, f32._bases, f32.ntotal = ex._chunks, ex._bases, ex.ntotal
s_f, i_f = f32.search_device(q, 5)
print("exact mode == float64 top-5:", (i_ex == best_i[:, :5]).all(dim=1).float().mean().item(), "rechecked", ex.last_rechecked)
print("fp32 kernel == float64 top-5:", (i_f == best_i[:, :5]).all(dim=1).float().mean().item())
bad = (~(i_f == best_i[:, :5]).all(dim=1)).nonzero().flatten()[:3]
for b in bad.tolist():
    print("query", b, "f64", best_i[b, :5].tolist(), [round(v, 6) for v in best_s[b, :5].tolist()])
    print("         fp32", i_f[b].tolist(), [round(v, 6) for v in s_f[b].tolist()])
    print("        exact", i_ex[b].tolist(), [round(v, 6) for v in s_ex[b].tolist()])
