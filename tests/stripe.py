"""One rank of a W-rank loss job replayed on a single GPU (test stand-in for the exchange step only: the kernels
are the product's).  Used by tests/test_gpu_loss.py and tools/stripe_step.py."""
import torch


class StripeComm:
    """One rank of a W-rank job, replayed on a single GPU: the gather returns the operands of the WHOLE batch
    (prepared row by row, so they equal what the ranks would have exchanged), the statistics all-reduce returns the
    sum the test accumulated over a first pass through every rank."""

    def __init__(self, ops, rank, world, feats_all, labels_all, stats_sum=None):
        self.ops, self.rank, self.world_size = ops, rank, world
        self.feats_all, self.labels_all = feats_all, labels_all
        self.stats_sum, self.stats_local = stats_sum, None

    def gather_tc16(self, feats, labels):
        xh, inv = [], []
        for f in self.feats_all:
            y16, iv = (None, None) if f is None else self.ops.rows_to_f16(f)
            xh.append(y16)
            inv.append(iv)
        return xh, inv, self.labels_all

    def all_gather_rows(self, t):
        raise AssertionError("the TC16 path gathers through gather_tc16")

    def all_reduce_sum(self, t):
        self.stats_local = t.clone()
        return t if self.stats_sum is None else self.stats_sum.clone()


def _local(feats_all, labels_all, rank, world):
    nl = labels_all.shape[0] // world
    sl = slice(rank * nl, (rank + 1) * nl)
    return [None if f is None else f[sl].clone() for f in feats_all], labels_all[sl].contiguous()


def first_pass(ops, world, feats_all, labels_all, scale=1 / 0.07):
    """Forward of every rank's stripe; returns the summed statistics vector (what the all-reduce would hand out)."""
    from bsc._lib import MODE_TC16
    from bsc.loss import _InfoNCEFn

    total = None
    for r in range(world):
        comm = StripeComm(ops, r, world, feats_all, labels_all)
        local, lab = _local(feats_all, labels_all, r, world)
        with torch.no_grad():
            _InfoNCEFn.apply(ops, comm, scale, MODE_TC16, 1.0, lab, *local)
        total = comm.stats_local if total is None else total + comm.stats_local
    return total


def rank_step(ops, rank, world, feats_all, labels_all, stats_sum, scale=1 / 0.07):
    """Forward + backward of one rank's stripe with the global statistics; returns (loss, leaves with .grad)."""
    from bsc._lib import MODE_TC16
    from bsc.loss import _InfoNCEFn

    comm = StripeComm(ops, rank, world, feats_all, labels_all, stats_sum)
    local, lab = _local(feats_all, labels_all, rank, world)
    leaves = [None if f is None else f.requires_grad_(True) for f in local]
    loss = _InfoNCEFn.apply(ops, comm, scale, MODE_TC16, 1.0, lab, *leaves)
    loss.backward()
    return loss.detach(), leaves
