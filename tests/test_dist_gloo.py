"""world_size-2 gloo test of the multi-rank host logic (row partition, all_gather, statistics all_reduce,
gradient gain) with the CPU stand-in ops; the reference's own 2-rank ClipLoss fixture is the expectation."""
import os

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

NAMES = ["image", "dna", "text"]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, feats, labels, gather_with_grad, ret):
    import sys

    for p in (ROOT, os.path.join(ROOT, "bioscan-clip_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist

    from bsc.loss import TorchDistComm, infonce_backward, infonce_forward
    from bsc._lib import MODE_F32
    from tests.fake_ops import FakeOps

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    n = labels.shape[0] // world
    sl = slice(rank * n, (rank + 1) * n)
    local = [None if f is None else f[sl].contiguous() for f in feats]
    ops, comm = FakeOps(), TorchDistComm(rank, world)
    loss, saved = infonce_forward(ops, comm, local, labels[sl].contiguous(), 1 / 0.07, MODE_F32)
    gain = float(world) if gather_with_grad else 1.0
    grads = infonce_backward(ops, saved, local, torch.ones(()), gain)
    ret[rank] = (float(loss), [None if g is None else g.numpy() for g in grads])
    dist.destroy_process_group()


def _run(feats, labels, gather_with_grad=False, port=29633):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, port, feats, labels, gather_with_grad, ret), nprocs=2, join=True)
    return ret


def test_two_ranks_match_reference_cliploss_fixture(golden_dir):
    z = np.load(os.path.join(golden_dir, "cliploss_w2.npz"))
    feats = [torch.from_numpy(z[f"x_{n}"]) for n in NAMES]
    labels = torch.from_numpy(z["labels"])
    ret = _run(feats, labels)
    for r in range(2):
        loss, grads = ret[r]
        assert abs(loss - float(z[f"loss_rank{r}"])) <= 1e-5 * abs(float(z[f"loss_rank{r}"]))
        for k, nm in enumerate(NAMES):
            ref = z[f"g_{nm}_rank{r}"]
            assert np.abs(grads[k] - ref).max() <= 1e-5 * np.abs(ref).max()


def test_two_ranks_missing_modality_and_gather_with_grad():
    from oracle import loss_oracle
    from tests.synth import loss_inputs

    feats, labels = loss_inputs(24, 16, (True, False, True), seed=2, normalised=False, labels="multi")
    ref_loss, ref_grads = loss_oracle.closed_form_grads(feats, labels)
    ret = _run(feats, labels, gather_with_grad=True, port=29634)
    for r in range(2):
        loss, grads = ret[r]
        assert abs(loss - float(ref_loss)) <= 1e-5 * abs(float(ref_loss))
        assert grads[1] is None
        for k in (0, 2):
            ref = 2.0 * ref_grads[k][r * 12:(r + 1) * 12].numpy()  # gather_with_grad: world_size x the slice
            assert np.abs(grads[k] - ref).max() <= 1e-5 * np.abs(ref).max()
