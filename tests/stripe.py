"""A W-rank loss job replayed on a single GPU (test stand-in for the exchange steps only: every kernel is the
product's, called through the C ABI).  Used by tests/test_gpu_loss.py and tools/stripe_step.py.

The replay runs the ranks one after the other, phase by phase, which is exactly the dependency order of the real job:
  1. forward of every rank's row stripe          -> the per-rank statistics records, handed out as the exchange would
  2. bsc_infonce_bwd_partials of every rank      -> in the scatter strategy every rank stores its partial gradients
                                                    into ALL ranks' exchange buffers (W tensors on the one device, handed
                                                    out as "peer" addresses)
  3. bsc_infonce_bwd_finish of every rank        -> fixed-order sum of the W sources + normalize backward
"""
import torch


class StripeComm:
    """One rank of a W-rank job: the gather returns the operands of the WHOLE batch (prepared row by row, so they equal
    what the ranks would have exchanged), the statistics all-reduce returns the sum accumulated over a first pass."""
    peer_memory = True

    def __init__(self, ops, rank, world, feats_all, labels_all, records=None, shared=None):
        self.ops, self.rank, self.world_size = ops, rank, world
        self.feats_all, self.labels_all = feats_all, labels_all
        self.records, self.stats_local = records, None
        self.shared = shared if shared is not None else {}
        self._x3_calls = 0

    def gather_tc16(self, feats, labels):
        if "gathered" not in self.shared:
            xh, inv = [], []
            for f in self.feats_all:
                y16, iv = (None, None) if f is None else self.ops.rows_to_f16(f)
                xh.append(y16)
                inv.append(iv)
            self.shared["gathered"] = (xh, inv)
        xh, inv = self.shared["gathered"]
        return {"xh": xh, "inv": inv, "labels": self.labels_all}

    def all_gather_rows(self, t):
        """BSC_MODE_TC16X3 gathers with plain collectives, in a fixed order: (hi | lo rows, inverse norms) of every
        present modality, then the labels.  The stand-in answers each call with the same quantity prepared from the
        WHOLE batch and checks that the caller's local part is the matching slice of it."""
        if "x3" not in self.shared:
            seq = []
            for f in self.feats_all:
                if f is not None:
                    seq.extend(self.ops.rows_to_f16x3(f))
            seq.append(self.labels_all)
            self.shared["x3"] = seq
        full = self.shared["x3"][self._x3_calls]
        self._x3_calls += 1
        nl = full.shape[0] // self.world_size
        assert torch.equal(full[self.rank * nl:(self.rank + 1) * nl], t), "gather order changed: update tests/stripe.py"
        return full

    def share_stats(self, record):
        """First pass (records is None): remember this rank's record.  Second pass: hand out all ranks' records, this
        rank's slot refreshed with the record just computed -- what the exchange over peer memory / all_gather yields."""
        self.stats_local = record.clone()
        if self.records is None:
            return record[None]
        out = self.records.clone()
        out[self.rank] = record
        return out

    def prepare_exchange(self, nbytes, device):
        return True


def _local(feats_all, labels_all, rank, world):
    nl = labels_all.shape[0] // world
    sl = slice(rank * nl, (rank + 1) * nl)
    return [None if f is None else f[sl].clone() for f in feats_all], labels_all[sl].contiguous()


def first_pass(ops, world, feats_all, labels_all, scale=1 / 0.07, shared=None, mode=None):
    """Forward of every rank's stripe; returns the [world, len] statistics records (what the exchange hands out)."""
    from bsc._lib import MODE_TC16
    from bsc.loss import infonce_forward

    MODE_TC16 = MODE_TC16 if mode is None else mode

    recs = []
    shared = shared if shared is not None else {}
    for r in range(world):
        comm = StripeComm(ops, r, world, feats_all, labels_all, shared=shared)
        local, lab = _local(feats_all, labels_all, r, world)
        infonce_forward(ops, comm, local, lab, scale, MODE_TC16, want_grad=False)
        recs.append(comm.stats_local)
    return torch.stack(recs)


def replay(ops, world, feats_all, labels_all, request, scale=1 / 0.07, grad_out=1.0, mode=None):
    """Whole job: returns (loss per rank, gradient per modality for ALL rows, resolved backward mode).
    mode: BSC_MODE_TC16 (default) or BSC_MODE_TC16X3."""
    from bsc._lib import BWD_SCATTER, MODE_TC16
    from bsc.loss import infonce_forward

    MODE_TC16 = MODE_TC16 if mode is None else mode
    shared = {}
    records = first_pass(ops, world, feats_all, labels_all, scale, shared, mode)
    nl = labels_all.shape[0] // world
    ranks = []
    for r in range(world):
        comm = StripeComm(ops, r, world, feats_all, labels_all, records, shared)
        local, lab = _local(feats_all, labels_all, r, world)
        loss, (prob, aux) = infonce_forward(ops, comm, local, lab, scale, MODE_TC16, True, request)
        ranks.append((loss, prob, aux, local))
    mode = ranks[0][1].bwd_mode
    dev = labels_all.device
    bufs, ptrs = [None] * world, None
    if mode == BWD_SCATTER:
        # poisoned: a slot nobody writes shows up as NaN in the result
        bufs = [torch.full((max(ranks[0][1].exchange_bytes // 4, 4),), float("nan"), dtype=torch.float32, device=dev)
                for _ in range(world)]
        ptrs = [b.data_ptr() for b in bufs]
    for loss, prob, aux, local in ranks:
        ops.infonce_bwd_partials(prob, aux, ptrs)
    g_out = torch.tensor(grad_out, dtype=torch.float32, device=dev)
    grads = [None if f is None else torch.empty_like(f) for f in feats_all]
    for r, (loss, prob, aux, local) in enumerate(ranks):
        gl = ops.infonce_bwd_finish(prob, aux, local, g_out, bufs[r])
        for g, l in zip(grads, gl):
            if g is not None:
                g[r * nl:(r + 1) * nl] = l
    return [float(x[0]) for x in ranks], grads, mode
