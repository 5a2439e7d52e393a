"""Seeded synthetic inputs."""
import numpy as np
import torch


def loss_inputs(n, d, present=(True, True, True), seed=0, normalised=True, labels="arange", shared=0.5,
                dtype=torch.float32, device="cpu"):
    g = torch.Generator().manual_seed(1000 + seed)
    z = torch.randn(n, d, generator=g)
    feats = []
    for m in range(3):
        if not present[m]:
            feats.append(None)
            continue
        x = z + shared * torch.randn(n, d, generator=g) if shared > 0 else torch.randn(n, d, generator=g)
        if normalised:
            x = torch.nn.functional.normalize(x, dim=1)
        else:
            x = x * (0.5 + 1.5 * torch.rand(n, 1, generator=g))
        feats.append(x.to(dtype).to(device))
    if labels == "arange":
        lab = torch.arange(n)
    elif labels == "multi":
        lab = torch.randint(0, max(2, n // 4), (n,), generator=g)
    elif labels == "same":
        lab = torch.zeros(n, dtype=torch.int64)
    else:
        raise ValueError(labels)
    return feats, lab.to(device)


def bank_inputs(nb, nq, d, seed=0, noise=0.3, n_dup=8, n_random=0):
    """Unit-norm bank, queries planted next to random bank rows, a few exact duplicate rows (tie -> lowest id)."""
    rng = np.random.default_rng(21 + seed)
    bank = rng.standard_normal((nb, d)).astype(np.float32)
    bank /= np.linalg.norm(bank, axis=1, keepdims=True)
    pi = rng.integers(0, nb, size=nq)
    for t in range(min(n_dup, nq)):  # duplicate the target of query t at a HIGHER index
        src = pi[t]
        dst = int(rng.integers(src + 1, nb)) if src + 1 < nb else src
        bank[dst] = bank[src]
    q = bank[pi] + noise * rng.standard_normal((nq, d)).astype(np.float32) / np.sqrt(d)
    if n_random:
        q[-n_random:] = rng.standard_normal((n_random, d)).astype(np.float32)
    q /= np.linalg.norm(q, axis=1, keepdims=True)
    return bank, q.astype(np.float32), pi
