"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's taxonomy-accuracy bookkeeping.

Restates, in plain Python, what scripts/inference_and_eval.py does with a search result:
  * map_hits_to_labels      the label loop of make_prediction              (inference_and_eval.py:422-433)
  * top_k_micro_accuracy    share of queries whose true label is in the first k hits, per level   (:448-464)
  * top_k_macro_accuracy    mean over the ground-truth classes of the per-class hit rate          (:467-511)
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may import this module.

Pin: the reference functions themselves cannot be imported here (inference_and_eval.py needs faiss, hydra, h5py,
plotly, umap at import time, none installable); tests/golden/make_golden.py therefore extracts the SOURCE of the three
functions from /root/reference with `ast`, executes exactly that source, and stores its outputs in
tests/golden/accuracy_small.json -- this restatement is checked against those vectors.
"""
from __future__ import annotations

from typing import Dict, List, Sequence

LEVELS = ["order", "family", "genus", "species"]  # inference_and_eval.py LEVELS


def map_hits_to_labels(indices, keys_label) -> List[Dict[str, list]]:
    """inference_and_eval.py:422-433: for every query, per level, the labels of its hits in rank order."""
    out = []
    for row in indices:
        out.append({level: [keys_label[int(i)][level] for i in row] for level in LEVELS})
    return out


def top_k_micro_accuracy(pred_list, gt_list, k_list: Sequence[int]):
    """inference_and_eval.py:448-464."""
    n = len(pred_list)
    acc = {}
    for k in k_list:
        acc.setdefault(k, {})
        for level in LEVELS:
            hits = sum(1 for pred, gt in zip(pred_list, gt_list) if gt[level] in pred[level][:k])
            acc[k][level] = hits * 1.0 / n
    return acc


def top_k_macro_accuracy(pred_list, gt_list, k_list: Sequence[int] = None):
    """inference_and_eval.py:467-511: classes are the ground-truth labels in order of first appearance; the per-class
    ratios are summed in that order (float64) and divided by the number of classes."""
    if k_list is None:
        k_list = [1, 3, 5]
    macro, per_class = {}, {}
    for k in k_list:
        macro[k], per_class[k] = {}, {}
        for level in LEVELS:
            hit_count: Dict[object, int] = {}
            seen_count: Dict[object, int] = {}
            for pred, gt in zip(pred_list, gt_list):
                label = gt[level]
                hit_count.setdefault(label, 0)
                seen_count.setdefault(label, 0)
                if label in pred[level][:k]:
                    hit_count[label] += 1
                seen_count[label] += 1
            total = 0
            per_class[k][level] = {}
            for label in seen_count:
                ratio = hit_count[label] * 1.0 / seen_count[label]
                total = total + ratio
                per_class[k][level][label] = ratio
            macro[k][level] = total / len(seen_count)
    return macro, per_class
