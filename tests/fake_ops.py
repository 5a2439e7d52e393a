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
        x = x.double()
        return x, 1.0 / x.norm(dim=1).clamp_min(1e-12)

    def label_stats(self, labels):
        t = (labels[None, :] == labels[:, None])
        r = t.sum(1).double()
        return r, t, None, r.max().reshape(1)

    def infonce_fwd(self, p):
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

    def infonce_loss(self, p, stats):
        n = p.n_global
        present = [m for m in range(3) if p.xh[m] is not None]
        m = len(present)
        tot = 0.0
        for a in present:
            for b in present:
                if a != b:
                    tot = tot + (p.r * torch.log(stats[pair_id(a, b) * n: pair_id(a, b) * n + n])).sum()
        return ((tot - 2.0 * stats[6 * n]) / (n * m * (m - 1))).float()

    def infonce_bwd(self, p, stats, xs, grad_out):
        n = p.n_global
        present = [m for m in range(3) if p.xh[m] is not None]
        m = len(present)
        c = 1.0 / (n * m * (m - 1))
        sl = slice(p.row0, p.row0 + p.n_local)
        t = p.rep.double()
        grads = [None, None, None]
        for x in present:
            ghat = torch.zeros(p.n_local, p.d, dtype=torch.float64)
            for y in present:
                if y == x:
                    continue
                e = torch.exp(p.scale * p.xh[x][sl] @ p.xh[y].T - p.scale)
                u = p.r[sl] / stats[pair_id(x, y) * n: pair_id(x, y) * n + n][sl]
                v = p.r / stats[pair_id(y, x) * n: pair_id(y, x) * n + n]
                g = e * (u[:, None] + v[None, :]) - 2 * t[sl]
                ghat += c * p.scale * g @ p.xh[y]
            xl = xs[x].double()
            nrm = xl.norm(dim=1, keepdim=True).clamp_min(1e-12)
            xh = xl / nrm
            grads[x] = (float(grad_out) * (ghat - xh * (xh * ghat).sum(1, keepdim=True)) / nrm).to(xs[x].dtype)
        return grads
