"""On-disk container of extracted embeddings with the reference's logical schema.

The reference keeps ``/seen|unseen|key/<feature_type>`` datasets in one HDF5 file plus a ``labels_*.json`` with the keys
``seen_gt_dict / unseen_gt_dict / key_gt_dict`` (scripts/inference_and_eval.py:804-831 load, :863-884 save).  h5py is
not part of this image, so the default container is a directory with the same tree, one ``.npy`` per dataset
(``<root>/<split>/<feature_type>.npy``); when h5py can be imported, a path ending in ``.hdf5`` / ``.h5`` reads and writes
the reference's file format itself.  ``stream_to_index`` feeds a stored bank to ``bsc.IndexFlatIP`` in pinned chunks
so a 5M x 768 bank never has to exist as one host array.
"""
from __future__ import annotations

import json
import os
from typing import Dict, Optional

import numpy as np

ALL_TYPE_OF_FEATURES_OF_KEY = [  # scripts/inference_and_eval.py:36-43
    "encoded_image_feature", "encoded_dna_feature", "encoded_language_feature", "averaged_feature",
    "concatenated_feature", "all_key_features",
]
SPLITS = ["seen", "unseen", "key"]
_LABEL_KEYS = {"seen": "seen_gt_dict", "unseen": "unseen_gt_dict", "key": "key_gt_dict"}


def _is_hdf5(path: str) -> bool:
    return path.endswith((".hdf5", ".h5"))


def _h5py():
    try:
        import h5py  # noqa: WPS433

        return h5py
    except ImportError as e:  # pragma: no cover - depends on the image
        raise ImportError("this path names an HDF5 file but h5py is not installed; use a directory container") from e


def _to_numpy(x):
    if x is None:
        return None
    if hasattr(x, "detach"):
        x = x.detach().cpu().numpy()
    return np.asarray(x)


def save_extracted_features(features_path: str, labels_path: str, seen_dict: dict, unseen_dict: dict,
                            keys_dict: dict) -> None:
    """scripts/inference_and_eval.py:863-884: every feature type present in a split becomes ``/<split>/<type>``; the
    label lists go to ``labels_path`` as JSON under the reference's three keys."""
    splits = dict(zip(SPLITS, [seen_dict, unseen_dict, keys_dict]))
    if _is_hdf5(features_path):
        h5 = _h5py()
        with h5.File(features_path, "w") as f:
            for name, split in splits.items():
                group = f.create_group(name)
                for ft in ALL_TYPE_OF_FEATURES_OF_KEY:
                    if split.get(ft) is not None:
                        group.create_dataset(ft, data=_to_numpy(split[ft]))
    else:
        for name, split in splits.items():
            os.makedirs(os.path.join(features_path, name), exist_ok=True)
            for ft in ALL_TYPE_OF_FEATURES_OF_KEY:
                if split.get(ft) is not None:
                    np.save(os.path.join(features_path, name, ft + ".npy"), _to_numpy(split[ft]))
    with open(labels_path, "w") as f:
        json.dump({_LABEL_KEYS[name]: split["label_list"] for name, split in splits.items()}, f, indent=4)


def load_extracted_features(features_path: str, labels_path: str, mmap: bool = False):
    """scripts/inference_and_eval.py:804-831 -> (seen_dict, unseen_dict, keys_dict) with ``label_list`` filled in and
    ``all_key_features_label`` = the key labels three times (:831).  ``mmap=True`` maps the .npy files instead of
    reading them (directory container only)."""
    out: Dict[str, dict] = {name: {} for name in SPLITS}
    if _is_hdf5(features_path):
        h5 = _h5py()
        with h5.File(features_path, "r") as f:
            for name in SPLITS:
                for ft in ALL_TYPE_OF_FEATURES_OF_KEY:
                    if ft in f[name].keys():
                        out[name][ft] = f[name][ft][:]
    else:
        for name in SPLITS:
            for ft in ALL_TYPE_OF_FEATURES_OF_KEY:
                p = os.path.join(features_path, name, ft + ".npy")
                if os.path.exists(p):
                    out[name][ft] = np.load(p, mmap_mode="r" if mmap else None)
    with open(labels_path) as f:
        total = json.load(f)
    for name in SPLITS:
        out[name]["label_list"] = total[_LABEL_KEYS[name]]
    out["key"]["all_key_features_label"] = total["key_gt_dict"] + total["key_gt_dict"] + total["key_gt_dict"]
    return out["seen"], out["unseen"], out["key"]


def stream_to_index(index, features, chunk_rows: int = 262144, normalise: bool = False, ops=None) -> int:
    """Append a stored bank (array or memory map, [n, d]) to a ``bsc.IndexFlatIP`` chunk by chunk through a pinned
    staging buffer.  ``normalise=True`` applies the sklearn rule of make_prediction (inference_and_eval.py:416) on the
    device before the rows are stored.  Returns the number of rows added."""
    import torch

    from ._lib import NORM_SKLEARN

    n = features.shape[0]
    if n == 0:
        return 0
    if ops is None and normalise:
        from .ops import CudaOps

        ops = CudaOps()
    dev = index.devices[0]
    stage = torch.empty((min(chunk_rows, n), features.shape[1]), dtype=torch.float32).pin_memory()
    for lo in range(0, n, chunk_rows):
        hi = min(n, lo + chunk_rows)
        view = stage[:hi - lo]
        view.copy_(torch.from_numpy(np.array(features[lo:hi], dtype=np.float32)))  # np.array: own writable copy
        rows = view.to(dev, non_blocking=True)
        if normalise:
            rows = ops.normalize(rows, torch.float32, NORM_SKLEARN)
        index.add(rows)
        torch.cuda.current_stream(dev).synchronize()  # the staging buffer is reused by the next chunk
    return n
