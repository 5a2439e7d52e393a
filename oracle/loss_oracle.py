"""CPU oracle for BIOSCAN-CLIP's multi-modal InfoNCE loss.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may import this file; the product package (``bioscan-clip_b200/bsc``) never does.

What it restates (reference @ 382d5ff, paths relative to /root/reference):

* ``construct_label_metrix``          bioscanclip/model/loss_func.py:18-21
* ``ContrastiveLoss.forward``         bioscanclip/model/loss_func.py:29-54
* ``gather_features`` (default path)  bioscanclip/model/loss_func.py:84-89
* ``ClipLoss.forward``                bioscanclip/model/loss_func.py:117-165
* soft-target ``nn.CrossEntropyLoss`` as constructed in scripts/train_cl.py:190

Parity pin: the reference has no tests or golden vectors for this path.  The oracle is pinned by
executing the reference module itself (it imports with torch alone) on seeded inputs in the build
container; ``tests/golden/make_golden.py`` wrote the resulting inputs/outputs to
``tests/golden/loss_*.npz`` and ``tests/test_oracle_golden.py`` checks this restatement against them.

The restatement is a closed form, not a transcription: for every ordered pair (a, b) of present
modalities the reference evaluates ``CE(s * A^ B^T, T)`` with un-normalised multi-hot targets
``T_ij = [label_i == label_j]`` and mean reduction, i.e.

    L_ab = 1/N * sum_i ( r_i * logsumexp_j S_ij  -  sum_j T_ij S_ij ),     r_i = sum_j T_ij

and the loop at loss_func.py:39-53 appends every ordered pair twice, so the mean over the list is
the mean over ordered pairs.
"""
from __future__ import annotations

from typing import Optional, Sequence

import torch
import torch.nn.functional as F

TOO_FEW = "Too less element for calculating the contrastive loss."  # loss_func.py:35,142


def label_matrix(labels: torch.Tensor, dtype=torch.float32) -> torch.Tensor:
    """T_ij = [label_i == label_j] (loss_func.py:18-21); rows are NOT normalised."""
    return (labels[None, :] == labels[:, None]).to(dtype)


def _present(features: Sequence[Optional[torch.Tensor]]):
    return [f for f in features if f is not None]  # loss_func.py:30-31


def pair_terms(a_hat: torch.Tensor, b_hat: torch.Tensor, t: torch.Tensor, scale: float) -> torch.Tensor:
    """Soft-label CE of s*A^B^T against T with mean reduction (loss_func.py:46,50)."""
    s = scale * a_hat @ b_hat.T
    r = t.sum(dim=1)
    return (r * torch.logsumexp(s, dim=1) - (t * s).sum(dim=1)).mean()


def contrastive_loss(features: Sequence[Optional[torch.Tensor]], labels: torch.Tensor,
                     logit_scale: float = 1 / 0.07, dtype: Optional[torch.dtype] = None) -> torch.Tensor:
    """ContrastiveLoss.forward (loss_func.py:29-54).

    ``features`` is (image, dna, text); ``None`` entries are dropped; fewer than two present raises
    ValueError with the reference's message.  ``dtype=None`` computes in the input dtype (the
    reference's behaviour: fp32, no autocast); pass torch.float64 for margin audits.
    """
    feats = _present(features)
    if len(feats) < 2:
        raise ValueError(TOO_FEW)
    if dtype is not None:
        feats = [f.to(dtype) for f in feats]
    t = label_matrix(labels, feats[0].dtype).to(feats[0].device)
    hats = [F.normalize(f, p=2, dim=1) for f in feats]  # loss_func.py:43-44 (eps 1e-12)
    total = 0.0
    count = 0
    for ia, a in enumerate(hats):
        for ib, b in enumerate(hats):
            if ia == ib:
                continue
            total = total + pair_terms(a, b, t, logit_scale)
            count += 1
    return total / count


def contrastive_loss_and_grads(features, labels, logit_scale=1 / 0.07, dtype=None):
    """Loss value plus d loss / d feature for every present modality (autograd on the restatement)."""
    leaves = [None if f is None else f.detach().clone().requires_grad_(True) for f in features]
    loss = contrastive_loss(leaves, labels, logit_scale, dtype)
    loss.backward()
    return loss.detach(), [None if l is None else l.grad for l in leaves]


def closed_form_grads(features, labels, logit_scale=1 / 0.07, dtype=torch.float64):
    """Analytic gradient used by the CUDA path, evaluated densely on the CPU for audits.

    dL/dS_ab[i,j] = (softmax_j(S_ab[i,:]) * r_i - T_ij) / (N * M(M-1))  (SURVEY.md section 0), then the
    chain through S = s*A^B^T and through F.normalize.
    """
    idx = [k for k, f in enumerate(features) if f is not None]
    xs = [features[k].to(dtype) for k in idx]
    m = len(xs)
    n = xs[0].shape[0]
    t = label_matrix(labels, dtype)
    r = t.sum(1)
    norms = [x.norm(dim=1, keepdim=True).clamp_min(1e-12) for x in xs]
    hats = [x / nn for x, nn in zip(xs, norms)]
    c = 1.0 / (n * m * (m - 1))
    ghat = [torch.zeros_like(x) for x in xs]
    loss = torch.zeros((), dtype=dtype)
    for a in range(m):
        for b in range(m):
            if a == b:
                continue
            s = logit_scale * hats[a] @ hats[b].T
            lse = torch.logsumexp(s, dim=1)
            loss = loss + c * (r * lse - (t * s).sum(1)).sum()
            g = c * (torch.softmax(s, dim=1) * r[:, None] - t)
            ghat[a] += logit_scale * g @ hats[b]
            ghat[b] += logit_scale * g.T @ hats[a]
    grads = [(g - h * (h * g).sum(1, keepdim=True)) / nn for g, h, nn in zip(ghat, hats, norms)]
    out = [None] * len(features)
    for k, g in zip(idx, grads):
        out[k] = g
    return loss, out


# ---------------------------------------------------------------------------------------------
# Distributed twin (ClipLoss) -- used by the gloo tests.
# ---------------------------------------------------------------------------------------------

def gather_rows_default(local: torch.Tensor, rank: int, world_size: int) -> torch.Tensor:
    """gather_features default branch (loss_func.py:84-89): no-grad all_gather, own slot re-inserted
    so only the local rows carry gradient.  Requires an initialised process group."""
    import torch.distributed as dist

    slots = [torch.zeros_like(local) for _ in range(world_size)]
    dist.all_gather(slots, local.detach())
    slots[rank] = local
    return torch.cat(slots, dim=0)


def clip_loss_rank(features, labels, logit_scale, rank: int, world_size: int, dtype=None):
    """ClipLoss.forward on one rank (loss_func.py:117-165): labels and every present modality are
    gathered, then the same pair loop runs on the GLOBAL batch on every rank."""
    import torch.distributed as dist

    lab_slots = [torch.zeros_like(labels) for _ in range(world_size)]
    dist.all_gather(lab_slots, labels)
    all_labels = torch.cat(lab_slots, dim=0)
    gathered = [None if f is None else (gather_rows_default(f, rank, world_size) if world_size > 1 else f)
                for f in features]
    return contrastive_loss(gathered, all_labels, logit_scale, dtype)


# ---------------------------------------------------------------------------------------------
# Block-wise evaluation for the headline sizes (N = 32768): never holds an N x N tensor.
# ---------------------------------------------------------------------------------------------

def blockwise_loss_and_grad_rows(features, labels, logit_scale=1 / 0.07, sample_rows=None, dtype=torch.float64,
                                 device=None, block=2048):
    """The closed form of ``closed_form_grads`` (loss_func.py:29-54 and its autograd) evaluated in row blocks.

    Pass 1: for every unordered pair (a, b) and every row block I, S = s * A^[I] B^^T (block x N) gives the row
    log-sum-exps of S_ab for I, a log-add-exp update of the column log-sum-exps (= the row log-sum-exps of S_ba)
    and the positive term sum_ij T_ij S_ij.  Pass 2: for ``sample_rows`` (1-D int64 tensor of row indices, the same
    rows for every modality; None = all rows) the gradient rows
        dA^_i = s * sum_j G_ij B^_j,   G_ij = c (r_i e^{S_ij - lse_ab[i]} + r_j e^{S_ij - lse_ba[j]} - 2 T_ij)
    chained through F.normalize.  Returns (loss, [grad rows per modality or None]) in ``dtype`` on the CPU.
    ``device`` may be a CUDA device: the matmuls are then plain torch fp64/fp32 GEMMs there -- test infrastructure
    on the same footing as running the reference module on that device.
    """
    idx = [k for k, f in enumerate(features) if f is not None]
    if len(idx) < 2:
        raise ValueError(TOO_FEW)
    dev = torch.device(device) if device is not None else features[idx[0]].device
    xs = [features[k].to(device=dev, dtype=dtype) for k in idx]
    lab = labels.to(dev)
    m, n = len(xs), xs[0].shape[0]
    norms = [x.norm(dim=1, keepdim=True).clamp_min(1e-12) for x in xs]
    hats = [x / nn for x, nn in zip(xs, norms)]
    # r_i = #{j: label_j == label_i} without the N x N matrix
    _, inverse, counts = torch.unique(lab, return_inverse=True, return_counts=True)
    r = counts[inverse].to(dtype)
    c = 1.0 / (n * m * (m - 1))
    neg_inf = float("-inf")
    lse = {}
    pos_total = torch.zeros((), dtype=dtype, device=dev)
    for a in range(m):
        for b in range(a + 1, m):
            row_lse = torch.empty(n, dtype=dtype, device=dev)
            col_lse = torch.full((n,), neg_inf, dtype=dtype, device=dev)
            for lo in range(0, n, block):
                hi = min(n, lo + block)
                s = logit_scale * hats[a][lo:hi] @ hats[b].T
                row_lse[lo:hi] = torch.logsumexp(s, dim=1)
                col_lse = torch.logaddexp(col_lse, torch.logsumexp(s, dim=0))
                t = (lab[lo:hi, None] == lab[None, :])
                pos_total = pos_total + (s * t).sum()
                del s, t
            lse[(a, b)] = row_lse
            lse[(b, a)] = col_lse
    loss = torch.zeros((), dtype=dtype, device=dev)
    for key, v in lse.items():
        loss = loss + (r * v).sum()
    loss = c * (loss - 2.0 * pos_total)  # every unordered pair's positive term appears in both directions

    rows = torch.arange(n, device=dev) if sample_rows is None else sample_rows.to(dev)
    out = [None] * len(features)
    for a in range(m):
        ghat = torch.zeros((rows.numel(), xs[a].shape[1]), dtype=dtype, device=dev)
        for b in range(m):
            if b == a:
                continue
            for lo in range(0, rows.numel(), block):
                sel = rows[lo:lo + block]
                s = logit_scale * hats[a][sel] @ hats[b].T
                g = r[sel, None] * torch.exp(s - lse[(a, b)][sel, None]) + r[None, :] * torch.exp(s - lse[(b, a)][None, :])
                g = g - 2.0 * (lab[sel, None] == lab[None, :]).to(dtype)
                ghat[lo:lo + block] += (c * logit_scale) * (g @ hats[b])
                del s, g
        h = hats[a][rows]
        out[idx[a]] = ((ghat - h * (h * ghat).sum(1, keepdim=True)) / norms[a][rows]).cpu()
    return loss.cpu(), out
