"""Drop-in for the reference's closest-embedding retrieval.

* ``IndexFlatIP(d)`` with ``add`` / ``search`` / ``ntotal`` / ``reset`` -- the faiss surface used at
  scripts/inference_and_eval.py:415-421, bioscanclip/epoch/eval_epoch.py:153-154,203 and
  bioscanclip/util/util.py:163-165.  ``search`` returns ``(float32[nq,k], int64[nq,k])`` host arrays, scores
  descending, equal scores ordered by lowest id, ``-1`` / ``-3.4028235e38`` padding when k > ntotal.
* ``make_prediction(query_feature, keys_feature, keys_label, with_similarity, with_indices, max_k)``
  -- scripts/inference_and_eval.py:414-445, same return forms.

The bank lives in HBM, split row-wise over ``devices`` at ``add`` time (one shard per GPU); a search scans
every shard on its own GPU and merges the per-shard top-k with ``bsc_topk_merge``.
``precision="exact"`` (the default) returns float32 results -- indices comparable bit for bit with a float32 CPU search,
which is what faiss ``IndexFlatIP`` computes -- at tensor-core speed: bf16 screening pass + exact re-scoring of the
candidates + a proof per query, unproven queries re-run exhaustively.  ``precision="fp32"`` is the exhaustive CUDA-core
float32 kernel itself (same results, the small-problem checker); ``precision="bf16"`` stores the bank in bfloat16 and
scans it with the tcgen05 kernel (bandwidth mode: bf16-rounded scores).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np
import torch

from ._lib import MODE_F32, MODE_TC16, NORM_SKLEARN, BscError

LEVELS = ["order", "family", "genus", "species"]


def _as_f32_tensor(x, device) -> torch.Tensor:
    if isinstance(x, torch.Tensor):
        return x.to(device=device, dtype=torch.float32).contiguous()
    arr = np.ascontiguousarray(np.asarray(x), dtype=np.float32)  # faiss takes float32 (inference_and_eval.py:416)
    return torch.from_numpy(arr).to(device)


class IndexFlatIP:
    def __init__(self, d: int, precision: str = "exact", devices: Optional[Sequence] = None, ops=None,
                 base_id: int = 0):
        if precision not in ("fp32", "bf16", "exact"):
            raise ValueError("precision must be 'fp32', 'bf16' or 'exact'")
        self.d = int(d)
        self.precision = precision
        if devices is None:
            devices = [torch.device("cuda", torch.cuda.current_device())] if torch.cuda.is_available() else []
        self.devices = [torch.device(dv) for dv in devices]
        if not self.devices:
            raise BscError("IndexFlatIP needs at least one CUDA device; there is no CPU path")
        if ops is None:
            from .ops import CudaOps

            ops = CudaOps()
        self._ops = ops
        self._mode = MODE_TC16 if precision == "bf16" else MODE_F32
        self._dtype = torch.bfloat16 if precision == "bf16" else torch.float32
        # "exact": float32 results (same as "fp32") found through a bf16 tensor-core screening pass over a bf16 copy of
        # the bank and exact re-scoring of the candidates; unproven queries fall back to the exhaustive float32 kernel
        self._screen: List[List[tuple]] = [[] for _ in self.devices]
        self.last_rechecked = 0
        self._base_id = int(base_id)
        self._chunks: List[List[torch.Tensor]] = [[] for _ in self.devices]  # per device: list of (base, rows)
        self._bases: List[List[int]] = [[] for _ in self.devices]
        self.ntotal = 0

    def reset(self) -> None:
        self._chunks = [[] for _ in self.devices]
        self._bases = [[] for _ in self.devices]
        self._screen = [[] for _ in self.devices]
        self.ntotal = 0

    def add(self, x) -> None:
        """Append rows (no normalisation, as faiss).  Rows are split evenly over the devices."""
        n = x.shape[0]
        if n == 0:
            return
        if x.shape[1] != self.d:
            raise ValueError(f"expected vectors of dimension {self.d}, got {x.shape[1]}")
        nd = len(self.devices)
        per = (n + nd - 1) // nd
        for k, dev in enumerate(self.devices):
            lo, hi = min(n, k * per), min(n, (k + 1) * per)
            if hi <= lo:
                continue
            rows = x[lo:hi]
            if isinstance(rows, torch.Tensor) and rows.dtype == self._dtype:
                part = rows.to(dev).contiguous()  # already in the bank's storage type: no float32 round trip
            else:
                part = _as_f32_tensor(rows, dev).to(self._dtype).contiguous()
            self._chunks[k].append(part)
            self._bases[k].append(self._base_id + self.ntotal + lo)
            if self.precision == "exact":
                max_norm = float(torch.linalg.vector_norm(part, dim=1).max().item())
                self._screen[k].append((part.to(torch.bfloat16).contiguous(), max_norm))
        self.ntotal += n

    def search_device(self, q, k: int):
        """Search returning device tensors (scores, ids) on the first device.  Every shard is scanned on its own GPU, on
        that GPU's current stream: all launches are queued before the first host read, so the shards run concurrently."""
        dev0 = self.devices[0]
        q0 = _as_f32_tensor(q, dev0)
        self.last_rechecked = 0
        if q0.shape[1] != self.d:
            raise ValueError(f"expected queries of dimension {self.d}, got {q0.shape[1]}")
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream(dev0))
        pending = []   # (device, queries on it, rows, base, launched-or-result)
        for kdev, dev in enumerate(self.devices):
            if not self._chunks[kdev]:
                continue
            with torch.cuda.device(dev):
                if dev != dev0:
                    torch.cuda.current_stream(dev).wait_event(ready)
                qd = q0.to(dev, non_blocking=True).to(self._dtype).contiguous()
                for ci, (base, rows) in enumerate(zip(self._bases[kdev], self._chunks[kdev])):
                    if self.precision == "exact":
                        rows16, max_norm = self._screen[kdev][ci]
                        launched = self._ops.bank_search_exact_launch(qd, rows, rows16, k, base, max_norm)
                        if launched is None:
                            launched = self._ops.bank_search(qd, rows, k, base, MODE_F32) + (None, None)
                            self.last_rechecked += qd.shape[0]
                        pending.append((dev, qd, rows, base, launched))
                    else:
                        pending.append((dev, qd, rows, base, self._ops.bank_search(qd, rows, k, base, self._mode)))
        cand_s, cand_i = [], []
        for dev, qd, rows, base, res in pending:
            with torch.cuda.device(dev):
                if self.precision == "exact" and res[2] is not None:
                    s, i, nbad = self._ops.bank_search_exact_finish(res, qd, rows, k, base)
                    self.last_rechecked += nbad
                else:
                    s, i = res[0], res[1]
                if dev != dev0:
                    done = torch.cuda.Event()
                    done.record(torch.cuda.current_stream(dev))
                    s, i = s.to(dev0, non_blocking=True), i.to(dev0, non_blocking=True)
                    torch.cuda.current_stream(dev0).wait_event(done)
            cand_s.append(s)
            cand_i.append(i)
        nq = q0.shape[0]
        if not cand_s:
            cs = torch.empty((0, nq, k), dtype=torch.float32, device=dev0)
            ci = torch.empty((0, nq, k), dtype=torch.int64, device=dev0)
        elif len(cand_s) == 1:
            return cand_s[0], cand_i[0]
        else:
            with torch.cuda.device(dev0):
                cs, ci = torch.stack(cand_s), torch.stack(cand_i)
        return self._ops.topk_merge(cs, ci)

    def search(self, q, k: int):
        s, i = self.search_device(q, k)
        return s.cpu().numpy(), i.cpu().numpy()


def merge_across_ranks(scores: torch.Tensor, ids: torch.Tensor, ops=None, group=None):
    """One process per GPU, the bank rows sharded over the ranks (every rank's index built with ``base_id`` = its shard's
    first global row), the queries replicated: this rank's ``[nq, k]`` result of ``IndexFlatIP.search_device`` -> the
    global ``[nq, k]`` on every rank.  ONE ``all_gather_into_tensor`` carries both lists (the score bits and the id of a
    candidate travel as an int64 pair), then ``bsc_topk_merge`` orders the world * k candidates of every query
    (score desc, id asc) -- the merge faiss does inside ``search`` when an index is sharded."""
    import torch.distributed as dist

    world = dist.get_world_size(group)
    if world == 1:
        return scores, ids
    if ops is None:
        from .ops import CudaOps

        ops = CudaOps()
    packed = torch.stack([scores.contiguous().view(torch.int32).to(torch.int64), ids.to(torch.int64)], dim=-1)
    nq = packed.shape[0]
    out = torch.empty((world * nq,) + tuple(packed.shape[1:]), dtype=torch.int64, device=packed.device)
    dist.all_gather_into_tensor(out, packed.contiguous(), group=group)
    out = out.view((world, nq) + tuple(packed.shape[1:]))
    cand_s = out[..., 0].contiguous().to(torch.int32).view(torch.float32)
    cand_i = out[..., 1].contiguous()
    return ops.topk_merge(cand_s, cand_i)


def _normalized_f32(ops, x, device) -> torch.Tensor:
    """sklearn ``normalize(x, norm='l2', axis=1).astype(np.float32)`` (inference_and_eval.py:416-417) on the device.
    float64 input (what inference_epoch.py:67 produces) is normalised IN float64 and rounded once, like the reference;
    float32 input is normalised in float32, as sklearn does for float32 arrays."""
    if isinstance(x, torch.Tensor):
        t = x.to(device)
        if t.dtype not in (torch.float32, torch.float64):
            t = t.to(torch.float32)
    else:
        arr = np.asarray(x)
        if arr.dtype not in (np.float32, np.float64):
            arr = arr.astype(np.float64)
        t = torch.from_numpy(np.ascontiguousarray(arr)).to(device)
    return ops.normalize(t.contiguous(), torch.float32, NORM_SKLEARN)


def make_prediction(query_feature, keys_feature, keys_label, with_similarity=False, with_indices=False, max_k=5,
                    precision: str = "exact", devices=None, ops=None):
    """scripts/inference_and_eval.py:414-445 on the GPU: L2-normalise both sides with sklearn's zero rule,
    float32, inner-product top-max_k, then map the hits to their taxonomy labels."""
    if ops is None:
        from .ops import CudaOps

        ops = CudaOps()
    index = IndexFlatIP(keys_feature.shape[-1], precision=precision, devices=devices, ops=ops)
    dev0 = index.devices[0]
    keys = _normalized_f32(ops, keys_feature, dev0)
    queries = _normalized_f32(ops, query_feature, dev0)
    index.add(keys)
    similarities, indices = index.search(queries, max_k)
    # label mapping of :422-433, one object-array gather per level instead of a Python loop over every hit
    # (bsc.TaxonomyEval does the accuracy bookkeeping on the device without building these lists at all)
    per_level = {}
    for level in LEVELS:
        names = np.empty(len(keys_label), dtype=object)
        names[:] = [kl[level] for kl in keys_label]
        per_level[level] = names[indices].tolist()  # a padding id -1 indexes from the end, as keys_label[-1] does there
    pred_list = [{level: per_level[level][q] for level in LEVELS} for q in range(indices.shape[0])]
    out = [pred_list]
    if with_similarity:
        out.append(similarities)
    if with_indices:
        out.append(indices)
    return out[0] if len(out) == 1 else out
