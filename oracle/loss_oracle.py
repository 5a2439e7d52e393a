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
