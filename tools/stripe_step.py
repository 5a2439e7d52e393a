"""Kernel times of ONE rank's share of the bench workload, replayed on one GPU (rank 0 of W at N = 32768): what a rank of
a W-GPU job spends in the library's forward and backward, without the exchange (the scatter backward's "peer" buffers
are W tensors on the one device, so its NVLink stores become HBM stores).
    python tools/stripe_step.py [world] [n] [steps] [scatter|blocks|recompute]"""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "bioscan-clip_b200"))
import torch  # noqa: E402

import bench  # noqa: E402
from bsc import _lib  # noqa: E402
from bsc.loss import _default_ops, infonce_forward  # noqa: E402
from tests.stripe import StripeComm, _local, first_pass  # noqa: E402

world = int(sys.argv[1]) if len(sys.argv) > 1 else 2
n = int(sys.argv[2]) if len(sys.argv) > 2 else 32768
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 5
request = _lib.BWD_NAMES[sys.argv[4]] if len(sys.argv) > 4 else _lib.BWD_SCATTER
torch.cuda.set_device(0)
dev = torch.device("cuda", 0)
feats, labels = bench.synth_loss_rows(torch, n, 0, n, dev, torch.bfloat16)
ops = _default_ops()
shared = {}
records = first_pass(ops, world, feats, labels, bench.SCALE, shared)
local, lab = _local(feats, labels, 0, world)
g_out = torch.ones((), device=dev)
bufs = None
_lib.set_timing(True)
f, b, g, w, h = [], [], [], [], []
for k in range(2 + steps):
    t0 = torch.cuda.Event(enable_timing=True)
    t1 = torch.cuda.Event(enable_timing=True)
    shared.pop("gathered", None)  # the stand-in gather (operand preparation of all N rows) is part of the step
    t0.record()
    c0 = time.perf_counter()
    comm = StripeComm(ops, 0, world, feats, labels, records, shared)
    loss, (prob, aux) = infonce_forward(ops, comm, local, lab, bench.SCALE, _lib.MODE_TC16, True, request)
    if prob.exchange_bytes and bufs is None:
        bufs = [torch.zeros(prob.exchange_bytes // 4, dtype=torch.float32, device=dev) for _ in range(world)]
    ops.infonce_bwd_partials(prob, aux, None if bufs is None else [t.data_ptr() for t in bufs])
    ops.infonce_bwd_finish(prob, aux, local, g_out, None if bufs is None else bufs[0])
    h1 = time.perf_counter() - c0  # host time to queue the step
    t1.record()
    torch.cuda.synchronize()
    if k >= 2:
        f.append(_lib.last_kernel_ms(0))
        b.append(_lib.last_kernel_ms(1))
        g.append(_lib.last_kernel_ms(3))
        w.append(t0.elapsed_time(t1))
        h.append(h1 * 1e3)
print("world %d n %d mode %d loss %.6f fwd_ms %.3f bwd_ms %.3f gemm_ms %.3f step_ms %.3f host_ms %.3f "
      "(step includes the stand-in gather)" % (world, n, _lib.last_bwd_mode(), float(loss), sum(f) / len(f),
                                                sum(b) / len(b), sum(g) / len(g), sum(w) / len(w), sum(h) / len(h)))
