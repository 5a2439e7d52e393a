"""CPU stand-in for bsc.ops.CudaOps, used ONLY by the gloo tests of the multi-rank host logic.

It evaluates, densely and in float64, exactly the quantities the C ABI is specified to produce
(include/bsc_hot.h): sum-exp statistics relative to the static maximum s, partial column sums over the local
rows, the positive-pair term, and the gradient of the local rows.  It lives under tests/ so the product
package cannot reach it.
"""
import torch


def pair_id(a, b):
    return a * 2 + (b - 1 if b > a else b)


class FakeOps:
    name = "fake-cpu"

    def normalize(self, x, out_dtype, zero_rule=0):
        x = x.double()
        n = x.norm(dim=1, keepdim=True)
        return (x / n.clamp_min(1e-12)).to(torch.float64)

    def rows_to_f16(self, x):
        """(operand rows, per-row scale) with cos_ij = <row_i, row_j> * scale_i * scale_j: the stand-in hands out
        the unit rows themselves and unit scales, so its formulas can ignore the scales."""
        x = x.double()
        return x / x.norm(dim=1, keepdim=True).clamp_min(1e-12), torch.ones(x.shape[0], dtype=torch.float64)

    def label_stats(self, labels):
        t = (labels[None, :] == labels[:, None])
        r = t.sum(1).double()
        return r, t, None, r.max().reshape(1)

    def infonce_fwd(self, p, exchange=None):
        n = p.n_global
        stats = torch.zeros(6 * n + 4, dtype=torch.float64)
        sl = slice(p.row0, p.row0 + p.n_local)
        t = p.rep.double()
        for a in range(3):
            for b in range(a + 1, 3):
                if p.xh[a] is None or p.xh[b] is None:
                    continue
                e = torch.exp(p.scale * p.xh[a][sl] @ p.xh[b].T - p.scale)
                stats[pair_id(a, b) * n + p.row0: pair_id(a, b) * n + p.row0 + p.n_local] = e.sum(1)
                stats[pair_id(b, a) * n: pair_id(b, a) * n + n] += e.sum(0)
        pos = 0.0  # sum over unordered pairs and local rows of sum_j T_ij (s*S_ij - s)
        for a in range(3):
            for b in range(a + 1, 3):
                if p.xh[a] is None or p.xh[b] is None:
                    continue
                pos = pos + ((p.scale * p.xh[a][sl] @ p.xh[b].T - p.scale) * t[sl]).sum()
        stats[6 * n] = pos
        return stats

    def infonce_loss(self, p, records):
        """bsc_infonce_loss: per-rank records added in rank order -> (loss, combined record, aux)."""
        stats = torch.zeros_like(records[0])
        for w in range(records.shape[0]):
            stats = stats + records[w]
        n = p.n_global
        present = [m for m in range(3) if p.xh[m] is not None]
        m = len(present)
        tot = 0.0
        for a in present:
            for b in present:
                if a != b:
                    tot = tot + (p.r * torch.log(stats[pair_id(a, b) * n: pair_id(a, b) * n + n])).sum()
        return ((tot - 2.0 * stats[6 * n]) / (n * m * (m - 1))).float(), stats, stats

    # ---- backward: the two phases of include/bsc_hot.h -------------------------------------------------------
    def plan(self, p, requested=-1, peer_memory=True):
        """bsc_infonce_plan: scatter when the ranks can reach each other's exchange buffers, blocks otherwise."""
        present = [m for m in range(3) if p.xh[m] is not None]
        np_ = len(present) * (len(present) - 1) // 2
        p.bwd_mode = requested if requested >= 0 else (4 if peer_memory else 3)
        p.exchange_bytes = p.world * np_ * p.n_local * p.d * 8 if p.bwd_mode == 4 else 0

    def _weights(self, p, stats, x, y):
        """G_xy[local rows, all columns] = e (u + v) - 2 T."""
        n = p.n_global
        sl = slice(p.row0, p.row0 + p.n_local)
        e = torch.exp(p.scale * p.xh[x][sl] @ p.xh[y].T - p.scale)
        u = p.r[sl] / stats[pair_id(x, y) * n: pair_id(x, y) * n + n][sl]
        v = p.r / stats[pair_id(y, x) * n: pair_id(y, x) * n + n]
        return e * (u[:, None] + v[None, :]) - 2 * p.rep.double()[sl]

    def infonce_bwd_partials(self, p, stats, exchange_ptrs=None):
        present = [m for m in range(3) if p.xh[m] is not None]
        sl = slice(p.row0, p.row0 + p.n_local)
        p.fake_local = {}
        if p.bwd_mode != 4:  # every ordered pair from this rank's own weight block: no exchange
            for x in present:
                for y in present:
                    if y != x:
                        p.fake_local[(x, y)] = self._weights(p, stats, x, y) @ p.xh[y]
            return
        # scatter: one weight block per unordered pair; dA for the local rows, the PARTIAL dB of every global row goes
        # to its owner's buffer (exchange_ptrs = outbox [dest rank][pair][n_local][d] of the test comm)
        outbox = exchange_ptrs
        q = 0
        for ia, a in enumerate(present):
            for b in present[ia + 1:]:
                g = self._weights(p, stats, a, b)
                p.fake_local[(a, b)] = g @ p.xh[b]
                part = g.T @ p.xh[a][sl]  # [n_global, d]
                outbox[:, q] = part.reshape(p.world, p.n_local, p.d)
                q += 1

    def infonce_bwd_finish(self, p, stats, xs, grad_out, own_exchange=None):
        n = p.n_global
        present = [m for m in range(3) if p.xh[m] is not None]
        m = len(present)
        c = 1.0 / (n * m * (m - 1))
        pair_index = {}
        for ia, a in enumerate(present):
            for b in present[ia + 1:]:
                pair_index[(a, b)] = len(pair_index)
        grads = [None, None, None]
        for x in present:
            ghat = torch.zeros(p.n_local, p.d, dtype=torch.float64)
            for y in present:
                if y == x:
                    continue
                if (x, y) in p.fake_local:
                    ghat += c * p.scale * p.fake_local[(x, y)]
                else:  # fixed-order sum over the source ranks
                    acc = torch.zeros(p.n_local, p.d, dtype=torch.float64)
                    for src in range(p.world):
                        acc += own_exchange[src, pair_index[(y, x)]]
                    ghat += c * p.scale * acc
            xl = xs[x].double()
            nrm = xl.norm(dim=1, keepdim=True).clamp_min(1e-12)
            xh = xl / nrm
            grads[x] = (float(grad_out) * (ghat - xh * (xh * ghat).sum(1, keepdim=True)) / nrm).to(xs[x].dtype)
        return grads
