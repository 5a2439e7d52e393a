"""How does the tensor-core accumulator round?  Scores of the bf16 bank scan are the raw fp32 accumulators of
tcgen05.mma kind::f16 (K = d in steps of 16); operands are exact bf16 values, so the fp64 dot product of the same values
is the exact answer.  Prints mean / std / extremes of (score - exact) / exact for self matches (all products positive)
and for planted neighbours (cos ~ 0.8), per width.
    python tools/acc_probe.py"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "bioscan-clip_b200"))
import numpy as np
import torch

import bsc

torch.manual_seed(0)
for d in (256, 768, 1536, 3072):
    nb, nq = 4096, 512
    bank = torch.nn.functional.normalize(torch.randn(nb, d), dim=1).bfloat16().float()
    for kind in ("self", "near", "far"):
        if kind == "self":
            q = bank[:nq].clone()
        elif kind == "near":
            q = torch.nn.functional.normalize(bank[:nq] + 0.5 * torch.randn(nq, d) / d ** 0.5, dim=1).bfloat16().float()
        else:
            q = torch.nn.functional.normalize(torch.randn(nq, d), dim=1).bfloat16().float()
        index = bsc.IndexFlatIP(d, precision="bf16")
        index.add(bank.numpy())
        s, i = index.search(q.numpy(), 1)
        exact = (q.double() * bank[torch.from_numpy(i[:, 0])].double()).sum(1).numpy()
        rel = (s[:, 0].astype(np.float64) - exact) / np.abs(exact)
        ulp = 2.0 ** -23
        print(f"d={d} {kind}: exact ~{np.mean(np.abs(exact)):.3f}  rel err mean {rel.mean():+.3e} std {rel.std():.2e} "
              f"min {rel.min():+.2e} max {rel.max():+.2e}   (mean / 2^-23 = {rel.mean() / ulp:+.2f}, per K=16 step "
              f"{rel.mean() / ulp / (d / 16):+.4f})")
