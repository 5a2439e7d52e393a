"""Taxonomy bookkeeping of a retrieval result on the GPU -- the step after the nearest-match search.

Drop-ins for scripts/inference_and_eval.py:
  * ``top_k_micro_accuracy(pred_list, gt_list, k_list)``            (:448-464), same return dict
  * ``top_k_macro_accuracy(pred_list, gt_list, k_list=None)``       (:467-511), same (macro, per_class) tuple
and the form that skips the Python label loop of make_prediction (:422-433) altogether:
  * ``TaxonomyEval(keys_label).evaluate(indices, gt_list, k_list)`` takes the int64 hit indices of
    ``IndexFlatIP.search`` (host array or device tensor) and returns (micro, macro, per_class).

The label strings of every level are interned to int32 ids on the host (ground-truth labels first, in order of first
appearance, which is the reference's class order); the counting runs in ``bsc_taxonomy_eval`` (integer atomics, exact);
the final ratios are formed in float64 in the reference's summation order, so the results compare equal (``==``) with
the reference's.  No CPU fallback for the counting: a missing library or device raises ``BscError``.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import _lib
from ._lib import BscError

LEVELS = ["order", "family", "genus", "species"]


class _Codec:
    """label -> id per level; ids are handed out in order of first appearance."""

    def __init__(self):
        self.maps: List[Dict[object, int]] = [dict() for _ in LEVELS]

    def encode(self, labels: Sequence[dict], grow: bool = True) -> np.ndarray:
        out = np.empty((len(labels), len(LEVELS)), dtype=np.int32)
        for li, level in enumerate(LEVELS):
            m = self.maps[li]
            col = out[:, li]
            for i, lab in enumerate(labels):
                v = lab[level]
                idx = m.get(v)
                if idx is None:
                    if grow:
                        idx = m[v] = len(m)
                    else:
                        idx = -1
                col[i] = idx
        return out

    def names(self, li: int) -> list:
        return list(self.maps[li].keys())


def _device(device) -> torch.device:
    if device is None:
        if not torch.cuda.is_available():
            raise BscError("taxonomy evaluation needs a CUDA device; there is no CPU path")
        return torch.device("cuda", torch.cuda.current_device())
    return torch.device(device)


def _run(indices_dev: torch.Tensor, key_ids: torch.Tensor, gt_ids: torch.Tensor, k_list: Sequence[int], n_classes,
         want_pred: bool = False):
    """-> micro[nk,levels] int64, pred_counts[nk,total] int32, gt_counts[total] int32, class_off, (pred_ids)"""
    dev = indices_dev.device
    _lib.require_device(dev)
    nq, k = indices_dev.shape
    levels = len(LEVELS)
    nk = len(k_list)
    off = np.zeros(levels, dtype=np.int64)
    off[1:] = np.cumsum(n_classes)[:-1]
    total = int(np.sum(n_classes))
    micro = torch.empty((nk, levels), dtype=torch.int64, device=dev)
    pred_counts = torch.empty((nk, max(total, 1)), dtype=torch.int32, device=dev)
    gt_counts = torch.empty(max(total, 1), dtype=torch.int32, device=dev)
    pred_ids = torch.empty((nq, k, levels), dtype=torch.int32, device=dev) if want_pred else None
    kl = (ctypes.c_int32 * nk)(*[int(v) for v in k_list])
    co = (ctypes.c_int64 * levels)(*[int(v) for v in off])
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().bsc_taxonomy_eval(
            indices_dev.data_ptr(), nq, k, key_ids.data_ptr(), key_ids.shape[0], gt_ids.data_ptr(), levels, kl, nk, co,
            max(total, 1), None if pred_ids is None else pred_ids.data_ptr(), None, micro.data_ptr(),
            pred_counts.data_ptr(), gt_counts.data_ptr(), _lib.stream_ptr(dev)))
    return micro, pred_counts, gt_counts, off, pred_ids


def _finish(micro, pred_counts, gt_counts, off, n_gt_classes, names, nq, k_list):
    micro = micro.cpu().numpy()
    pc = pred_counts.cpu().numpy()
    gc = gt_counts.cpu().numpy()
    micro_acc, macro_acc, per_class = {}, {}, {}
    for ki, k in enumerate(k_list):
        micro_acc.setdefault(k, {})
        macro_acc[k], per_class[k] = {}, {}
        for li, level in enumerate(LEVELS):
            micro_acc[k][level] = int(micro[ki, li]) * 1.0 / nq
            g = n_gt_classes[li]
            lo = int(off[li])
            # ids 0..g-1 of a level are exactly the ground-truth classes in order of first appearance
            ratios = pc[ki, lo:lo + g].astype(np.float64) / gc[lo:lo + g].astype(np.float64)
            total = float(np.cumsum(ratios)[-1]) if g else 0.0  # cumsum adds left to right like the reference loop
            macro_acc[k][level] = total / g
            per_class[k][level] = dict(zip(names[li][:g], ratios.tolist()))
    return micro_acc, macro_acc, per_class


class TaxonomyEval:
    """Holds the bank's taxonomy labels as device-resident id columns."""

    def __init__(self, keys_label: Sequence[dict], device=None):
        self.device = _device(device)
        self.keys_label = keys_label
        self._gt_cache = None

    def evaluate(self, indices, gt_list: Sequence[dict], k_list: Optional[Sequence[int]] = None,
                 with_pred_ids: bool = False):
        if k_list is None:
            k_list = [1, 3, 5]
        k_list = list(k_list)
        codec = _Codec()
        gt_np = codec.encode(gt_list)                      # ground truth first: reference class order
        n_gt = [len(m) for m in codec.maps]
        key_np = codec.encode(self.keys_label)             # unseen key labels get later ids (never counted)
        n_cls = [len(m) for m in codec.maps]
        if isinstance(indices, torch.Tensor):
            idx = indices.to(device=self.device, dtype=torch.int64).contiguous()
        else:
            idx = torch.from_numpy(np.ascontiguousarray(np.asarray(indices), dtype=np.int64)).to(self.device)
        if idx.shape[0] != len(gt_list):
            raise ValueError("one ground-truth label per query expected")
        key_ids = torch.from_numpy(key_np).to(self.device)
        gt_ids = torch.from_numpy(gt_np).to(self.device)
        micro, pc, gc, off, pred_ids = _run(idx, key_ids, gt_ids, k_list, n_cls, with_pred_ids)
        names = [codec.names(li) for li in range(len(LEVELS))]
        out = _finish(micro, pc, gc, off, n_gt, names, len(gt_list), k_list)
        return out + (pred_ids,) if with_pred_ids else out


def _from_pred_list(pred_list, gt_list, k_list, device):
    """Reference-signature entry: the predicted label strings are interned and fed to the same kernel as a bank whose
    row q*k+t is hit t of query q."""
    dev = _device(device)
    nq = len(pred_list)
    kmax = max(k_list)
    codec = _Codec()
    gt_np = codec.encode(gt_list)
    n_gt = [len(m) for m in codec.maps]
    flat = np.full((nq * kmax, len(LEVELS)), -1, dtype=np.int32)
    for li, level in enumerate(LEVELS):
        m = codec.maps[li]
        for q, pred in enumerate(pred_list):
            for t, v in enumerate(pred[level][:kmax]):
                idx = m.get(v)
                if idx is None:
                    idx = m[v] = len(m)
                flat[q * kmax + t, li] = idx
    n_cls = [len(m) for m in codec.maps]
    idx = torch.arange(nq * kmax, dtype=torch.int64, device=dev).view(nq, kmax)
    micro, pc, gc, off, _ = _run(idx, torch.from_numpy(flat).to(dev), torch.from_numpy(gt_np).to(dev), k_list, n_cls)
    names = [codec.names(li) for li in range(len(LEVELS))]
    return _finish(micro, pc, gc, off, n_gt, names, nq, k_list)


def top_k_micro_accuracy(pred_list, gt_list, k_list=None, device=None):
    """scripts/inference_and_eval.py:448-464 (k_list is required there too: None raises TypeError)."""
    if k_list is None:
        raise TypeError("'NoneType' object is not iterable")
    return _from_pred_list(pred_list, gt_list, list(k_list), device)[0]


def top_k_macro_accuracy(pred_list, gt_list, k_list=None, device=None):
    """scripts/inference_and_eval.py:467-511."""
    if k_list is None:
        k_list = [1, 3, 5]
    _, macro, per_class = _from_pred_list(pred_list, gt_list, list(k_list), device)
    return macro, per_class
