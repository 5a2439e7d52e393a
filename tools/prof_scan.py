"""Bank scan of the bench workload (bf16 bank, 768-d): a few 32-query passes and one bulk pass; the command ncu wraps.

    python tools/prof_scan.py [bank_rows] [bulk_queries]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "bioscan-clip_b200"))
import torch  # noqa: E402

import bsc  # noqa: E402
from bsc import _lib  # noqa: E402

nb = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
nq = int(sys.argv[2]) if len(sys.argv) > 2 else 8192
torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(21)
bank = torch.empty((nb, 768), dtype=torch.bfloat16, device=dev)
for lo in range(0, nb, 262144):
    hi = min(nb, lo + 262144)
    bank[lo:hi] = torch.nn.functional.normalize(torch.randn(hi - lo, 768, generator=g, device=dev), dim=1)
q = torch.nn.functional.normalize(torch.randn(nq, 768, generator=g, device=dev), dim=1)
index = bsc.IndexFlatIP(768, precision="bf16", devices=[dev])
index.add(bank)
_lib.set_timing(True)
for _ in range(3):
    s, i = index.search_device(q[:32].contiguous(), 5)
torch.cuda.synchronize()
print("stream32 scan kernel ms", _lib.last_kernel_ms(2), "GB/s", nb * 768 * 2 / (_lib.last_kernel_ms(2) * 1e-3) / 1e9)
s, i = index.search_device(q, 5)
torch.cuda.synchronize()
print("bulk scan kernel ms", _lib.last_kernel_ms(2), "TFLOP/s", 2.0 * nq * nb * 768 / (_lib.last_kernel_ms(2) * 1e-3) / 1e12)
