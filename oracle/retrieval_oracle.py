"""CPU oracle for BIOSCAN-CLIP's closest-embedding taxonomic retrieval.  TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / ``--impl reference``
legs may import this file; the product package never does.

What it restates (reference @ 382d5ff, paths relative to /root/reference):

* ``make_prediction``                scripts/inference_and_eval.py:414-445
* the ``IndexFlatIP.add/search`` use bioscanclip/epoch/eval_epoch.py:153-154,203 and
                                     bioscanclip/util/util.py:159-169

Third-party arithmetic that is NOT under /root/reference:

* ``faiss.IndexFlatIP`` (requirements.txt:22, ``faiss-gpu==1.7.2``; not installed, not vendored).
  Published algorithm restated here: exact brute-force inner product in float32, the k largest per
  query returned in descending score order with int64 ids; when k exceeds the bank size the tail is
  padded with id -1 and score -inf-like sentinel (faiss uses -3.4028235e38 for IP).
  faiss' order among EQUAL scores is an implementation detail; this oracle (and the CUDA path) fix it
  to "lowest index first" as BASELINE.json's north_star demands.
* ``sklearn.preprocessing.normalize(x, norm="l2", axis=1)`` (requirements.txt:4): rows divided by
  their L2 norm; rows whose norm is below 10*eps (all-zero rows included) are left unchanged.  Pinned against scikit-learn 1.9 in the build
  container by ``tests/golden/make_golden.py`` -> ``tests/golden/retrieval_normalize.npz``.

PARITY UNPINNED for the search itself: the reference ships no test or golden vector for it and
faiss cannot be executed here.  The pins that exist are the sklearn fixture above and an internal
cross-check (blocked top-k vs a full stable argsort) in ``tests/test_oracle_golden.py``.
"""
from __future__ import annotations

import numpy as np

LEVELS = ["order", "family", "genus", "species"]  # scripts/inference_and_eval.py (module constant)
FAISS_NEG = np.float32(-3.4028234663852886e38)


def l2_normalize_rows(x: np.ndarray) -> np.ndarray:
    """sklearn ``normalize(x, 'l2', axis=1)``: zero rows stay zero (inference_and_eval.py:416-417).
    Computed in the input dtype (float64 when fed the reference's list-built arrays)."""
    x = np.asarray(x)
    if x.dtype not in (np.float32, np.float64):
        x = x.astype(np.float64)
    norms = np.sqrt(np.einsum("ij,ij->i", x, x))
    # sklearn's _handle_zeros_in_scale: a norm below 10*eps(dtype) counts as zero and the row is left as is
    norms = np.where(norms < 10 * np.finfo(x.dtype).eps, 1, norms).astype(x.dtype)
    return x / norms[:, None]


def search_ip(keys: np.ndarray, queries: np.ndarray, k: int, block: int = 65536):
    """IndexFlatIP.add(keys); .search(queries, k) -> (float32[nq,k], int64[nq,k]).

    Exact float32 inner products, descending score, ties -> lowest index.  Blocked over the bank so
    the nq x nb score matrix is never materialised.
    """
    keys = np.ascontiguousarray(keys, dtype=np.float32)
    queries = np.ascontiguousarray(queries, dtype=np.float32)
    nq, nb = queries.shape[0], keys.shape[0]
    best_s = np.full((nq, k), FAISS_NEG, dtype=np.float32)
    best_i = np.full((nq, k), -1, dtype=np.int64)
    for lo in range(0, nb, block):
        hi = min(nb, lo + block)
        s = queries @ keys[lo:hi].T  # float32 sgemm
        cand_s = np.concatenate([best_s, s], axis=1)
        cand_i = np.concatenate([best_i, np.broadcast_to(np.arange(lo, hi, dtype=np.int64), s.shape)], axis=1)
        # sort by (-score, index): padding ids (-1) carry the sentinel score so they sort last
        key_i = np.where(cand_i < 0, np.iinfo(np.int64).max, cand_i)
        order = np.lexsort((key_i, -cand_s.astype(np.float64)), axis=1)[:, :k]
        best_s = np.take_along_axis(cand_s, order, axis=1)
        best_i = np.take_along_axis(cand_i, order, axis=1)
    return best_s, best_i


def search_ip_full_sort(keys: np.ndarray, queries: np.ndarray, k: int):
    """Cross-check: one dense score matrix + stable argsort (small sizes only)."""
    keys = np.ascontiguousarray(keys, dtype=np.float32)
    queries = np.ascontiguousarray(queries, dtype=np.float32)
    s = queries @ keys.T
    order = np.argsort(-s, axis=1, kind="stable")[:, :k]
    return np.take_along_axis(s, order, axis=1), order.astype(np.int64)


def make_prediction(query_feature, keys_feature, keys_label, with_similarity=False, with_indices=False, max_k=5):
    """scripts/inference_and_eval.py:414-445: normalise both sides, float32, IP search at max_k, then
    map every hit to its 4-level taxonomy labels.  Return forms follow :435-445."""
    keys = l2_normalize_rows(keys_feature).astype(np.float32)
    queries = l2_normalize_rows(query_feature).astype(np.float32)
    similarities, indices = search_ip(keys, queries, max_k)
    pred_list = []
    for row in indices:
        pred_list.append({level: [keys_label[int(i)][level] for i in row] for level in LEVELS})
    out = [pred_list]
    if with_similarity:
        out.append(similarities)
    if with_indices:
        out.append(indices)
    return out[0] if len(out) == 1 else out


def score_margins_fp64(keys, queries, indices):
    """fp64 audit helper: for each query the gap between the k-th returned score and the best
    non-returned score, to tell a genuine mismatch from a float32 near-tie."""
    keys64 = np.asarray(keys, dtype=np.float64)
    q64 = np.asarray(queries, dtype=np.float64)
    s = q64 @ keys64.T
    nq = s.shape[0]
    kth = s[np.arange(nq), indices[:, -1]]
    masked = s.copy()
    masked[np.arange(nq)[:, None], indices] = -np.inf
    return kth - masked.max(axis=1)


def search_ip_topk_torch(keys, queries, k: int, block: int = 262144, threads: int = 0):
    """The same search as search_ip with the selection done by torch.topk per bank block and a running merge --
    how a CPU IndexFlatIP spends its time (blocked sgemm + k-selection, inference_and_eval.py:418-421).  Used as the
    timed CPU baseline of the retrieval half of the metric; scores equal search_ip's, ids
    agree wherever the k-th and (k+1)-th scores differ (torch.topk does not promise the lowest id among ties)."""
    import torch

    if threads:
        torch.set_num_threads(threads)
    kt = torch.as_tensor(np.ascontiguousarray(keys, dtype=np.float32))
    qt = torch.as_tensor(np.ascontiguousarray(queries, dtype=np.float32))
    nq, nb = qt.shape[0], kt.shape[0]
    best_s = torch.full((nq, k), float(FAISS_NEG), dtype=torch.float32)
    best_i = torch.full((nq, k), -1, dtype=torch.int64)
    for lo in range(0, nb, block):
        hi = min(nb, lo + block)
        s = qt @ kt[lo:hi].T
        kk = min(k, hi - lo)
        ts, ti = torch.topk(s, kk, dim=1)
        cs = torch.cat([best_s, ts], dim=1)
        ci = torch.cat([best_i, ti + lo], dim=1)
        ms, mi = torch.topk(cs, k, dim=1)
        best_s, best_i = ms, torch.gather(ci, 1, mi)
    return best_s.numpy(), best_i.numpy()


def search_ip_blocked_topk(keys, queries, k: int, slack: int = 8, block: int = 262144, threads: int = 0):
    """search_ip for large banks: per bank block a float32 sgemm and ``torch.topk(k + slack)`` pick candidates, the
    final order (score desc, id asc) is decided by one lexsort over the few hundred candidates of a query -- the same
    result as search_ip unless more than k + slack rows of ONE block tie exactly at a query's k-th score.  Used by
    bench.py as the checker of the 5M-row search and as its timed CPU baseline (blocked sgemm + k-selection is how a
    CPU IndexFlatIP spends its time, inference_and_eval.py:418-421)."""
    import torch

    if threads:
        torch.set_num_threads(threads)
    kt = torch.as_tensor(np.ascontiguousarray(keys, dtype=np.float32)) if not hasattr(keys, "numpy") else keys
    qt = torch.as_tensor(np.ascontiguousarray(queries, dtype=np.float32))
    nq, nb = qt.shape[0], kt.shape[0]
    cs, ci = [], []
    for lo in range(0, nb, block):
        hi = min(nb, lo + block)
        s = qt @ kt[lo:hi].T
        kk = min(k + slack, hi - lo)
        ts, ti = torch.topk(s, kk, dim=1)
        cs.append(ts)
        ci.append(ti + lo)
    cand_s = torch.cat(cs, dim=1).numpy()
    cand_i = torch.cat(ci, dim=1).numpy()
    order = np.lexsort((cand_i, -cand_s.astype(np.float64)), axis=1)[:, :k]
    out_s = np.take_along_axis(cand_s, order, axis=1)
    out_i = np.take_along_axis(cand_i, order, axis=1)
    if out_s.shape[1] < k:  # fewer keys than k: faiss pads
        pad = k - out_s.shape[1]
        out_s = np.concatenate([out_s, np.full((nq, pad), FAISS_NEG, np.float32)], axis=1)
        out_i = np.concatenate([out_i, np.full((nq, pad), -1, np.int64)], axis=1)
    return out_s.astype(np.float32), out_i.astype(np.int64)
