"""Time of the fused gather (bsc_gather_push + symmetric-memory barrier) on real ranks, per-peer stores vs NVSwitch
multicast, against NCCL all_gather_into_tensor of the same bytes.
    python -m torch.distributed.run --nproc-per-node W --master-addr 127.0.0.1 tools/gather_probe.py [n_global]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "bioscan-clip_b200"))
import torch
import torch.distributed as dist

import bench
from bsc.loss import SymmMemComm

rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 32768
nl = n // world
feats, labels = bench.synth_loss_rows(torch, n, rank * nl, nl, dev, torch.bfloat16)


def timed(fn, iters=20):
    for _ in range(3):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    import time

    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    t0 = time.perf_counter()
    for _ in range(iters):
        fn()
    host = (time.perf_counter() - t0) / iters * 1e6
    b.record()
    b.synchronize()
    return round(a.elapsed_time(b) / iters * 1e3, 1), round(host, 1)  # (device us, host us to queue)


out = {}
for name, env in (("multicast", "1"), ("per-peer", "0")):
    os.environ["BSC_GATHER_MULTICAST"] = env
    comm = SymmMemComm(rank, world, strict=True)
    out[name] = timed(lambda: comm.gather_tc16(feats, labels))
    st = next(iter(comm._gather.values()))
    out[name + "_mcast_ptr"] = st["mcast"][0] != 0
# the kernel alone, no barrier, no Python around it: queue 20 launches through the C ABI directly
import ctypes

from bsc import _lib

st = next(iter(comm._gather.values()))
x = (ctypes.c_void_p * 3)(*[f.data_ptr() for f in feats])
ldx = (ctypes.c_int64 * 3)(*[f.stride(0) for f in feats])
lab = labels.to(torch.int64).contiguous()
for name, mc in (("kernel_only_per_peer", None), ("kernel_only_multicast", [h.multicast_ptr for h in st["handles"]][0] or None)):
    def push(mc=mc):
        _lib.check(_lib.lib().bsc_gather_push(x, _lib.dtype_code(feats[0].dtype), ldx, lab.data_ptr(), nl, rank * nl, n, 768,
                                              st["peers"][0], mc, world, _lib.stream_ptr(dev)))
    out[name] = timed(push)
out["barrier_only"] = timed(lambda: st["handles"][0].barrier(channel=0))
y = [f.to(torch.float16) for f in feats]
big = [torch.empty((n, 768), dtype=torch.float16, device=dev) for _ in range(3)]


def nccl():
    for src, dst in zip(y, big):
        dist.all_gather_into_tensor(dst, src)


out["nccl_3x_all_gather"] = timed(nccl)
if rank == 0:
    print("gather of 3 x [%d x 768] fp16 over %d ranks (us per call, back to back): %s" % (nl, world, out), flush=True)
dist.barrier()
dist.destroy_process_group()
