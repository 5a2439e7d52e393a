"""Does the ordered gather push (side stream) run BESIDE the main stream's kernels?  One GPU, world = 1 (the "peer" is a
plain device buffer): time of the label statistics and of a forward sweep alone and with the push in flight.
    python tools/overlap_probe.py [n]"""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "bioscan-clip_b200"))
import torch

import bench
from bsc import _lib
from bsc.ops import CudaOps

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
feats, labels = bench.synth_loss_rows(torch, n, 0, n, dev, torch.bfloat16)
L = _lib.lib()
offx, offi = (ctypes.c_size_t * 3)(), (ctypes.c_size_t * 3)()
offl, offs, offf, total = (ctypes.c_size_t(0) for _ in range(4))
_lib.check(L.bsc_gather_layout(n, 768, 1, offx, offi, ctypes.byref(offl), ctypes.byref(offs), ctypes.byref(offf), ctypes.byref(total)))
buf = torch.zeros(int(total.value), dtype=torch.uint8, device=dev)
peers = (ctypes.c_void_p * 1)(buf.data_ptr())
x = (ctypes.c_void_p * 3)(*[f.data_ptr() for f in feats])
ldx = (ctypes.c_int64 * 3)(*[f.stride(0) for f in feats])
side = torch.cuda.Stream(dev)
ops = CudaOps()
use_main = torch.cuda.Stream(dev) if os.environ.get("PROBE_MAIN_STREAM") == "1" else None


EPOCH = [0]


def push(stream, reps):
    for _ in range(reps):
        EPOCH[0] += 1
        _lib.check(L.bsc_gather_push(x, _lib.BSC_BF16, ldx, None, n, 0, n, 768, peers, None, 1, 0, 1, EPOCH[0],
                                     stream.cuda_stream))


def timed(fn, with_push):
    out = []
    for _ in range(5):
        torch.cuda.synchronize()
        if with_push:
            push(side, 6)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        out.append(a.elapsed_time(b) * 1e3)
    return sorted(out)[len(out) // 2]


def run():
    ctx = torch.cuda.stream(use_main) if use_main is not None else torch.cuda.stream(torch.cuda.current_stream())
    with ctx:
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(side)
        push(side, 1)
        b.record(side)
        torch.cuda.synchronize()
        print("push alone: %.1f us" % (a.elapsed_time(b) * 1e3))
        print("label_stats alone %.1f us, beside the push %.1f us" % (timed(lambda: ops.label_stats(labels), False),
                                                                      timed(lambda: ops.label_stats(labels), True)))
        y = torch.empty(1 << 26, dtype=torch.float32, device=dev)
        print("fill 256 MB alone %.1f us, beside the push %.1f us" % (timed(lambda: y.fill_(1.0), False),
                                                                      timed(lambda: y.fill_(1.0), True)))


run()
