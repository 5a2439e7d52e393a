"""world_size-2 gloo test of the multi-rank host logic (row partition, all_gather, statistics all_reduce,
gradient gain) with the CPU stand-in ops; the reference's own 2-rank ClipLoss fixture is the expectation."""
import os

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

NAMES = ["image", "dna", "text"]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


class GlooScatterComm:
    """Test stand-in for SymmMemComm on CPU: the 'peer memory' of the scatter backward is an outbox that the barrier
    delivers with an all_gather (rank r keeps block [r] of every rank's outbox, indexed by source rank)."""
    peer_memory = True

    def __init__(self, base):
        self.base, self.rank, self.world_size = base, base.rank, base.world_size
        self.all_gather_rows, self.share_stats = base.all_gather_rows, base.share_stats

    def prepare_exchange(self, nbytes, device):
        self.shape = None
        return True

    def exchange(self, nbytes, device):
        import torch

        self.outbox = torch.zeros(nbytes // 8, dtype=torch.float64)
        self.inbox = torch.zeros(nbytes // 8, dtype=torch.float64)
        return self.inbox, self.outbox

    def exchange_barrier(self, device):
        import torch
        import torch.distributed as dist

        boxes = [torch.zeros_like(self.outbox) for _ in range(self.world_size)]
        dist.all_gather(boxes, self.outbox)
        w = self.world_size
        per = self.outbox.numel() // w
        for src in range(w):
            self.inbox.view(w, per)[src] = boxes[src].view(w, per)[self.rank]


class _ShapedOps:
    """Gives the flat exchange tensors of GlooScatterComm the [rank][pair][n_local][d] shape FakeOps works on."""

    def __init__(self, ops):
        self.ops = ops

    def __getattr__(self, name):
        return getattr(self.ops, name)

    def _shape(self, p, t):
        return t.view(p.world, -1, p.n_local, p.d)

    def infonce_bwd_partials(self, p, stats, exchange_ptrs=None):
        return self.ops.infonce_bwd_partials(p, stats, None if exchange_ptrs is None else self._shape(p, exchange_ptrs))

    def infonce_bwd_finish(self, p, stats, xs, grad_out, own_exchange=None):
        return self.ops.infonce_bwd_finish(p, stats, xs, grad_out,
                                           None if own_exchange is None else self._shape(p, own_exchange))


def _worker(rank, world, port, feats, labels, gather_with_grad, ret, scatter=False):
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
    if scatter:
        ops, comm = _ShapedOps(ops), GlooScatterComm(comm)
    from bsc._lib import MODE_TC16

    # the TC16 host path (rows_to_f16 + gathers of rows and inverse norms) when the scatter backward is exercised
    mode = MODE_TC16 if scatter else MODE_F32
    loss, saved = infonce_forward(ops, comm, local, labels[sl].contiguous(), 1 / 0.07, mode)
    assert saved[0].bwd_mode == (4 if scatter else 3)
    gain = float(world) if gather_with_grad else 1.0
    grads = infonce_backward(ops, comm, saved, local, torch.ones(()), gain)
    ret[rank] = (float(loss), [None if g is None else g.numpy() for g in grads])
    dist.destroy_process_group()


def _run(feats, labels, gather_with_grad=False, port=29633, scatter=False):
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_worker, args=(2, port, feats, labels, gather_with_grad, ret, scatter), nprocs=2, join=True)
    return ret


@pytest.mark.parametrize("scatter", [False, True])
def test_two_ranks_match_reference_cliploss_fixture(golden_dir, scatter):
    """scatter=False: every rank forms its gradient from its own weight blocks (no backward exchange); scatter=True:
    the host protocol of the reduce-scatter backward (plan -> exchange buffers -> partials -> barrier -> finish).  Both
    against the reference's own 2-rank ClipLoss run."""
    z = np.load(os.path.join(golden_dir, "cliploss_w2.npz"))
    feats = [torch.from_numpy(z[f"x_{n}"]) for n in NAMES]
    labels = torch.from_numpy(z["labels"])
    ret = _run(feats, labels, scatter=scatter, port=29633 + (2 if scatter else 0))
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
    ret = _run(feats, labels, gather_with_grad=True, port=29634, scatter=True)
    for r in range(2):
        loss, grads = ret[r]
        assert abs(loss - float(ref_loss)) <= 1e-5 * abs(float(ref_loss))
        assert grads[1] is None
        for k in (0, 2):
            ref = 2.0 * ref_grads[k][r * 12:(r + 1) * 12].numpy()  # gather_with_grad: world_size x the slice
            assert np.abs(grads[k] - ref).max() <= 1e-5 * np.abs(ref).max()


class _TorchMerge:
    """CPU stand-in for CudaOps.topk_merge: [parts, nq, k] candidate lists -> [nq, k] by (score desc, id asc)."""

    def topk_merge(self, cand_s, cand_i):
        parts, nq, k = cand_s.shape
        s = cand_s.permute(1, 0, 2).reshape(nq, parts * k).numpy()
        i = cand_i.permute(1, 0, 2).reshape(nq, parts * k).numpy()
        out_s, out_i = np.empty((nq, k), np.float32), np.empty((nq, k), np.int64)
        for q in range(nq):
            order = np.lexsort((i[q], -s[q].astype(np.float64)))[:k]
            out_s[q], out_i[q] = s[q][order], i[q][order]
        return torch.from_numpy(out_s), torch.from_numpy(out_i)


def _merge_worker(rank, world, port, bank, queries, ret):
    import sys

    for p in (ROOT, os.path.join(ROOT, "bioscan-clip_b200")):
        if p not in sys.path:
            sys.path.insert(0, p)
    import torch.distributed as dist

    from bsc.retrieval import merge_across_ranks

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    nb = bank.shape[0] // world
    shard = bank[rank * nb:(rank + 1) * nb]
    sc = queries @ shard.T  # this rank's shard, ids offset by its first row; negative scores included
    s, i = torch.sort(sc, dim=1, descending=True, stable=True)
    s, i = s[:, :5].contiguous(), (i[:, :5] + rank * nb).contiguous()
    gs, gi = merge_across_ranks(s, i, ops=_TorchMerge())
    ret[rank] = (gs.numpy(), gi.numpy())
    dist.destroy_process_group()


def test_two_ranks_merge_of_sharded_search_results():
    """bsc.merge_across_ranks: the per-rank top-k lists travel in ONE all_gather (score bits + id packed in int64
    pairs) and come out as the top-k of the whole bank on every rank, negative scores and duplicate rows (ties ->
    lowest id) included.  The merge kernel itself is replaced by a numpy sort here (its GPU test is test_gpu_retrieval)."""
    g = torch.Generator().manual_seed(5)
    bank = torch.randn(64, 16, generator=g)
    bank[40] = bank[7]  # an exact duplicate on the other shard
    queries = torch.cat([bank[[7, 3, 50]] + 0.01 * torch.randn(3, 16, generator=g), -bank[[1]]])
    full = queries @ bank.T
    want_i = np.stack([np.lexsort((np.arange(64), -full[q].double().numpy()))[:5] for q in range(4)])
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_merge_worker, args=(2, 29641, bank, queries, ret), nprocs=2, join=True)
    for r in range(2):
        gs, gi = ret[r]
        np.testing.assert_array_equal(gi, want_i)
        np.testing.assert_allclose(gs, np.take_along_axis(full.numpy(), want_i, axis=1), rtol=0, atol=1e-6)
