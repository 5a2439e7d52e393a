"""32-query bank scan (the HBM-bound mode S of SURVEY 8d) over several shard sizes: kernel time by CUDA events with an
L2 flush between launches; the fit T = c + bytes / BW separates the fixed cost of a launch from the streaming rate.
    python tools/scan_probe.py [max_rows]          (BSC_LIB_PATH selects the build)
Also printed: GPU time per search of 20 searches queued back to back (scan + merge of the candidate lists, the host
running ahead; every shard is larger than L2, so no flush is needed between them)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "bioscan-clip_b200"))
import torch

import bsc
from bsc import _lib

torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(21)
nb_max = int(sys.argv[1]) if len(sys.argv) > 1 else 5_000_000
bank = torch.empty((nb_max, 768), dtype=torch.bfloat16, device=dev)
for lo in range(0, nb_max, 262144):
    hi = min(nb_max, lo + 262144)
    bank[lo:hi] = torch.nn.functional.normalize(torch.randn(hi - lo, 768, generator=g, device=dev), dim=1)
q = torch.nn.functional.normalize(torch.randn(32, 768, generator=g, device=dev), dim=1)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
_lib.set_timing(True)
for nb in (156_250, 312_500, 625_000, 1_250_000, 2_500_000, 5_000_000):
    if nb > nb_max:
        break
    index = bsc.IndexFlatIP(768, precision="bf16", devices=[dev])
    index.add(bank[:nb])
    ks, ws = [], []
    for it in range(8):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        s, i = index.search_device(q, 5)
        b.record()
        torch.cuda.synchronize()
        if it >= 3:
            ks.append(_lib.last_kernel_ms(2))
            ws.append(a.elapsed_time(b))
    k = sorted(ks)[len(ks) // 2]
    w = sorted(ws)[len(ws) // 2]
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for it in range(20):
        s, i = index.search_device(q, 5)
    b.record()
    torch.cuda.synchronize()
    bb = a.elapsed_time(b) / 20
    print(f"rows {nb:8d}  {nb * 1536 / 1e9:6.3f} GB  scan kernel {k * 1e3:7.1f} us = {nb * 1536 / k / 1e6:7.1f} GB/s   "
          f"whole search {w * 1e3:7.1f} us   back to back {bb * 1e3:7.1f} us per search")
