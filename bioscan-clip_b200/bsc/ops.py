"""Tensor-level wrappers over the C ABI: one method per entry point of include/bsc_hot.h.

``CudaOps`` is the only implementation the package ships.  The orchestration in ``loss.py`` takes the ops
object as an argument so that the multi-rank host logic can be exercised on CPU (gloo) by the tests with a
stand-in defined under ``tests/`` -- the product modules always construct ``CudaOps``.
"""
from __future__ import annotations

import ctypes
import os
from ctypes import byref, c_int, c_size_t, c_void_p
from dataclasses import dataclass, field
from typing import List, Optional

import torch

from . import _lib
from ._lib import MODE_F32, MODE_TC16, TC_MODES, NORM_SKLEARN, NORM_TORCH, InfoNceDesc, check, lib, ptr, stream_ptr


@dataclass
class Problem:
    """One InfoNCE evaluation: the gathered, normalised global batch and this rank's row range."""
    # 3 entries or None: F32 mode -> [n_global, d] float32 L2-normalised rows;
    # TC16 mode -> [n_global, d] float16 rows (the caller's values times a power of two) plus their inverse norms in ``inv``
    # TC16X3 mode -> [n_global, 2 d] float16 rows (hi | lo parts) plus ``inv``; ``d`` stays the embedding width
    xh: List[Optional[torch.Tensor]]
    n_global: int
    n_local: int
    row0: int
    d: int
    mode: int
    scale: float
    r: torch.Tensor = None      # [n_global] float32   label-matrix row sums
    rep: torch.Tensor = None    # [n_global] int32     class representative
    nxt: torch.Tensor = None    # [n_global] int32     next member of the class
    rmax: torch.Tensor = None   # [1] float32
    inv: List[Optional[torch.Tensor]] = None  # TC16: [n_global] float32 1 / max(||row||, 1e-12) per modality
    all_labels: torch.Tensor = None
    world: int = 1
    rank: int = 0
    # resolved by CudaOps.plan (bsc_infonce_plan)
    bwd_mode: int = None
    fwd_workspace_bytes: int = 0
    bwd_workspace_bytes: int = 0
    exchange_bytes: int = 0
    bwd_workspace: torch.Tensor = None
    _desc: InfoNceDesc = field(default=None, repr=False)

    def rebind(self, xh, inv, all_labels):
        """Point the problem at private copies of its operands (the symmetric gather buffer is about to be reused)."""
        self.xh, self.inv, self.all_labels = xh, inv, all_labels
        self._desc = None


def requested_bwd_mode(default: str = "auto") -> int:
    """The backward strategy the caller asked for: argument of the loss module, overridden by BSC_BWD_MODE
    (auto | recompute | blocks | scatter).  Read on the host once per forward; the library never reads it."""
    import os

    name = os.environ.get("BSC_BWD_MODE", default).strip().lower()
    if name not in _lib.BWD_NAMES:
        raise ValueError(f"backward mode must be one of {sorted(_lib.BWD_NAMES)} (got {name!r})")
    return _lib.BWD_NAMES[name]


def memory_budget(device) -> int:
    """Bytes the backward may claim: what the driver reports free plus what torch's caching allocator holds unused,
    minus a margin; BSC_BWD_MEM_GB caps it (the stored weight blocks are the one N^2-sized buffer of the design)."""
    import os

    free, _ = torch.cuda.mem_get_info(device)
    cached = torch.cuda.memory_reserved(device) - torch.cuda.memory_allocated(device)
    budget = int(0.9 * (free + max(cached, 0)))
    cap = os.environ.get("BSC_BWD_MEM_GB")
    if cap:
        budget = min(budget, int(float(cap) * (1 << 30)))
    return max(budget, 1)


class CudaOps:
    name = "cuda"

    def __init__(self):
        self._plans = {}
        self._ws = {}

    # ---- K1 -------------------------------------------------------------------------------------
    def normalize(self, x: torch.Tensor, out_dtype: torch.dtype, zero_rule: int = NORM_TORCH,
                  out: torch.Tensor = None) -> torch.Tensor:
        """Rows of x scaled to unit L2 norm (F.normalize / sklearn rule), written as out_dtype (into ``out``, a
        contiguous [n, d] tensor of that dtype, when given -- e.g. a slice of a preallocated embedding bank)."""
        _lib.require_device(x.device)
        assert x.dim() == 2
        if x.stride(1) != 1:
            x = x.contiguous()
        n, d = x.shape
        if out is not None:
            if out.shape != (n, d) or out.dtype != out_dtype or not out.is_contiguous() or out.device != x.device:
                raise _lib.BscError("normalize: out must be a contiguous [n, d] tensor of out_dtype on x's device")
            y = out
        else:
            y = torch.empty((n, d), dtype=out_dtype, device=x.device)
        y16 = y if out_dtype in (torch.float16, torch.bfloat16) else None
        y32 = y if out_dtype == torch.float32 else None
        if y16 is None and y32 is None:
            raise _lib.BscError(f"normalize: unsupported output dtype {out_dtype}")
        with torch.cuda.device(x.device):
            check(lib().bsc_l2norm_fwd(x.data_ptr(), _lib.dtype_code(x.dtype), n, d, x.stride(0), ptr(y16),
                                       _lib.dtype_code(out_dtype) if y16 is not None else _lib.BSC_F16, ptr(y32),
                                       0, zero_rule, stream_ptr(x.device)))
        return y

    def rows_to_f16(self, x: torch.Tensor):
        """TC16 operands: (x * 2^k_row as float16 -- exact for 16-bit inputs --, 2^-k_row / max(||x_row||, 1e-12))."""
        _lib.require_device(x.device)
        assert x.dim() == 2
        if x.stride(1) != 1:
            x = x.contiguous()
        n, d = x.shape
        y = torch.empty((n, d), dtype=torch.float16, device=x.device)
        inv = torch.empty(n, dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            check(lib().bsc_rows_to_f16(x.data_ptr(), _lib.dtype_code(x.dtype), n, d, x.stride(0), y.data_ptr(),
                                        inv.data_ptr(), stream_ptr(x.device)))
        return y, inv

    def rows_to_f16x3(self, x: torch.Tensor):
        """BSC_MODE_TC16X3 operands: ([n, 2 d] float16 = hi | lo parts of x * 2^k_row, 2^-k_row / max(||x_row||, 1e-12))."""
        _lib.require_device(x.device)
        assert x.dim() == 2
        if x.stride(1) != 1:
            x = x.contiguous()
        n, d = x.shape
        hl = torch.empty((n, 2 * d), dtype=torch.float16, device=x.device)
        inv = torch.empty(n, dtype=torch.float32, device=x.device)
        with torch.cuda.device(x.device):
            check(lib().bsc_rows_to_f16x3(x.data_ptr(), _lib.dtype_code(x.dtype), n, d, x.stride(0), hl.data_ptr(),
                                          inv.data_ptr(), stream_ptr(x.device)))
        return hl, inv

    # ---- K2 -------------------------------------------------------------------------------------
    def label_stats(self, labels: torch.Tensor):
        _lib.require_device(labels.device)
        labels = labels.to(torch.int64).contiguous()
        n = labels.numel()
        dev = labels.device
        r = torch.empty(n, dtype=torch.float32, device=dev)
        rep = torch.empty(n, dtype=torch.int32, device=dev)
        nxt = torch.empty(n, dtype=torch.int32, device=dev)
        rmax = torch.empty(1, dtype=torch.float32, device=dev)
        nbytes = c_size_t(0)
        check(lib().bsc_label_stats_workspace_bytes(n, byref(nbytes)))
        ws = torch.empty(int(nbytes.value), dtype=torch.uint8, device=dev)
        with torch.cuda.device(dev):
            check(lib().bsc_label_stats(labels.data_ptr(), n, r.data_ptr(), rep.data_ptr(), nxt.data_ptr(),
                                        rmax.data_ptr(), ws.data_ptr(), ws.numel(), stream_ptr(dev)))
        return r, rep, nxt, rmax

    # ---- InfoNCE --------------------------------------------------------------------------------
    def _desc(self, p: Problem) -> InfoNceDesc:
        if p._desc is None:
            d = InfoNceDesc()
            d.n_global, d.n_local, d.row0 = p.n_global, p.n_local, p.row0
            d.d, d.mode, d.logit_scale = p.d, p.mode, float(p.scale)
            first = next(t for t in p.xh if t is not None)
            d.x16_dtype = _lib.dtype_code(first.dtype) if p.mode in TC_MODES else 0
            for m in range(3):
                d.xh[m] = ptr(p.xh[m]) or None
                d.inv[m] = (ptr(p.inv[m]) if p.inv is not None else 0) or None
            d.r, d.rep, d.next, d.rmax = p.r.data_ptr(), p.rep.data_ptr(), p.nxt.data_ptr(), p.rmax.data_ptr()
            d.bwd_mode = p.bwd_mode if p.bwd_mode is not None else _lib.BWD_RECOMPUTE
            d.world, d.rank = p.world, p.rank
            p._desc = d
        return p._desc

    def _device(self, p: Problem) -> torch.device:
        return next(t for t in p.xh if t is not None).device

    def stats_len(self, n_global: int) -> int:
        return int(lib().bsc_infonce_stats_len(n_global))

    def plan(self, p: Problem, requested: int = _lib.BWD_AUTO, peer_memory: bool = True) -> None:
        """Resolve the backward strategy (bsc_infonce_plan) and record the buffer sizes in ``p``.  The verdict depends
        on the shape of the problem only, so it is asked for once per shape (querying the free device memory costs a
        driver call per step otherwise); ``forget_plans`` drops the cache, e.g. after an out-of-memory error."""
        dev = self._device(p)
        _lib.require_device(dev)
        key = (p.n_global, p.n_local, p.row0, p.d, p.mode, tuple(t is not None for t in p.xh), p.world, p.rank,
               requested, bool(peer_memory), dev, os.environ.get("BSC_BWD_MEM_GB"))
        hit = self._plans.get(key)
        if hit is None:
            out = _lib.InfoNcePlan()
            p._desc = None
            budget = memory_budget(dev) if (requested == _lib.BWD_AUTO and p.mode in TC_MODES) else 0
            check(lib().bsc_infonce_plan(byref(self._desc(p)), requested, 1 if peer_memory else 0, budget, byref(out)))
            hit = (int(out.bwd_mode), int(out.fwd_workspace_bytes), int(out.bwd_workspace_bytes),
                   int(out.exchange_bytes))
            if len(self._plans) > 64:
                self._plans.clear()
            self._plans[key] = hit
        p.bwd_mode, p.fwd_workspace_bytes, p.bwd_workspace_bytes, p.exchange_bytes = hit
        p._desc = None

    def forget_plans(self) -> None:
        self._plans.clear()

    def infonce_fwd(self, p: Problem, exchange: dict = None) -> torch.Tensor:
        """Forward sweep -> this rank's statistics record.  ``exchange`` (peer memory, see bsc_infonce_exchange):
        {"stats_peers": [addresses], "stats_multicast": address or 0, "arrive": address or 0, "arrive_target": int}."""
        dev = self._device(p)
        _lib.require_device(dev)
        if p.bwd_mode is None:
            self.plan(p)
        ws = torch.empty(max(p.fwd_workspace_bytes, 256), dtype=torch.uint8, device=dev)  # lives for the forward only
        stats = torch.empty(self.stats_len(p.n_global), dtype=torch.float32, device=dev)
        ex = None
        if exchange:
            ex = _lib.InfoNceExchange()
            peers = exchange.get("stats_peers")
            keep = (c_void_p * len(peers))(*peers) if peers else None
            ex.stats_peers = keep
            ex.stats_multicast = exchange.get("stats_multicast") or None
            ex.arrive = exchange.get("arrive") or None
            ex.arrive_target = int(exchange.get("arrive_target") or 0)
        with torch.cuda.device(dev):
            check(lib().bsc_infonce_fwd(byref(self._desc(p)), stats.data_ptr(), byref(ex) if ex is not None else None,
                                        ws.data_ptr(), ws.numel(), stream_ptr(dev)))
        return stats

    def infonce_loss(self, p: Problem, records: torch.Tensor):
        """records: [n_records, stats_len] per-rank statistics records (this rank's own as [1, stats_len] when all rows
        are local).  Returns (loss, combined record, aux = the backward's factors)."""
        dev = self._device(p)
        if records.dim() == 1:
            records = records[None]
        records = records.contiguous()
        loss = torch.empty((), dtype=torch.float32, device=dev)
        stats = torch.empty(self.stats_len(p.n_global), dtype=torch.float32, device=dev)
        aux = torch.empty(int(lib().bsc_infonce_aux_len(p.n_global)), dtype=torch.float32, device=dev)
        key = ("loss_scratch", dev, torch.cuda.current_stream(dev).cuda_stream)
        scratch = self._ws.get(key)
        if scratch is None:  # zeroed once; every call leaves it zeroed (ticket counter + block partial sums)
            scratch = self._ws[key] = torch.zeros(int(lib().bsc_infonce_loss_scratch_bytes()), dtype=torch.uint8,
                                                  device=dev)
        with torch.cuda.device(dev):
            check(lib().bsc_infonce_loss(byref(self._desc(p)), records.data_ptr(), records.shape[0], stats.data_ptr(),
                                         aux.data_ptr(), loss.data_ptr(), scratch.data_ptr(), stream_ptr(dev)))
        return loss, stats, aux

    def infonce_bwd_partials(self, p: Problem, aux: torch.Tensor, exchange_ptrs=None) -> None:
        """Phase 1 of the backward: raw gradient sums into the (lazily allocated) backward workspace and, in the
        scatter strategy, into the owners' exchange buffers (``exchange_ptrs``: one device address per rank)."""
        dev = self._device(p)
        if p.bwd_workspace is None:
            p.bwd_workspace = torch.empty(max(p.bwd_workspace_bytes, 256), dtype=torch.uint8, device=dev)
        ex = None
        if p.mode in TC_MODES and p.bwd_mode == _lib.BWD_SCATTER:
            if exchange_ptrs is None or len(exchange_ptrs) != p.world:
                raise _lib.BscError("the scatter backward needs one exchange buffer address per rank")
            ex = (c_void_p * p.world)(*exchange_ptrs)
        with torch.cuda.device(dev):
            check(lib().bsc_infonce_bwd_partials(byref(self._desc(p)), aux.data_ptr(), ex, p.bwd_workspace.data_ptr(),
                                                 p.bwd_workspace.numel(), stream_ptr(dev)))

    def infonce_bwd_finish(self, p: Problem, aux: torch.Tensor, xs, grad_out: torch.Tensor,
                           own_exchange: torch.Tensor = None):
        """Phase 2: xs = this rank's original rows per modality (None when absent; rows may be strided, elements must
        be contiguous).  Returns dense gradients of the same shapes and dtype."""
        dev = self._device(p)
        grad_out = grad_out.to(device=dev, dtype=torch.float32).reshape(1).contiguous()
        xs = [None if x is None else (x if x.stride(1) == 1 else x.contiguous()) for x in xs]
        present = [x for x in xs if x is not None]
        if any(x.dtype != present[0].dtype for x in present):
            raise _lib.BscError("all modalities must share one dtype (float32, bfloat16 or float16)")
        grads = [None if x is None else torch.empty(x.shape, dtype=x.dtype, device=dev) for x in xs]
        xp = (c_void_p * 3)(*[ptr(x) or None for x in xs])
        gp = (c_void_p * 3)(*[ptr(g) or None for g in grads])
        ldx = (ctypes.c_int64 * 3)(*[0 if x is None else x.stride(0) for x in xs])
        with torch.cuda.device(dev):
            check(lib().bsc_infonce_bwd_finish(byref(self._desc(p)), aux.data_ptr(), xp,
                                               _lib.dtype_code(present[0].dtype), ldx, grad_out.data_ptr(), gp,
                                               ptr(own_exchange) or None, p.bwd_workspace.data_ptr(),
                                               p.bwd_workspace.numel(), stream_ptr(dev)))
        return grads

    # ---- retrieval ------------------------------------------------------------------------------
    def _scratch(self, key, nbytes: int, device) -> torch.Tensor:
        """Workspace kept between searches (one per device and role): a search of the same shape allocates nothing."""
        buf = self._ws.get((key, device))
        if buf is None or buf.numel() < nbytes:
            buf = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=device)
            self._ws[(key, device)] = buf
        return buf

    def bank_search(self, queries: torch.Tensor, bank: torch.Tensor, k: int, base_id: int, mode: int):
        dev = bank.device
        _lib.require_device(dev)
        nq, d = queries.shape
        nb = bank.shape[0]
        nbytes = c_size_t(0)
        check(lib().bsc_bank_search_workspace_bytes(nq, nb, d, k, mode, byref(nbytes)))
        ws = self._scratch("search", int(nbytes.value), dev)
        scores = torch.empty((nq, k), dtype=torch.float32, device=dev)
        ids = torch.empty((nq, k), dtype=torch.int64, device=dev)
        with torch.cuda.device(dev):
            check(lib().bsc_bank_search(queries.data_ptr(), nq, bank.data_ptr(), nb, d, k, base_id, mode,
                                        scores.data_ptr(), ids.data_ptr(), ws.data_ptr(), ws.numel(),
                                        stream_ptr(dev)))
        return scores, ids

    def bank_search_exact_launch(self, queries32: torch.Tensor, bank32: torch.Tensor, bank16: torch.Tensor, k: int,
                                 base_id: int, max_row_norm: float, min_parts: int = 16):
        """Exact float32 top-k, device side only: bf16 screening pass + float32 re-scoring + per-query proof.  Returns
        (scores, ids, unsure flags, n_unsure) without synchronising; ``bank_search_exact_finish`` re-runs the unproven
        queries.  None when the shape is outside the screening kernel's range (the caller searches exhaustively)."""
        dev = bank32.device
        _lib.require_device(dev)
        nq, d = queries32.shape
        nb = bank32.shape[0]
        if k > 8 or d % 64 != 0 or nb == 0 or nq == 0:
            return None
        parts = c_int(0)
        check(lib().bsc_bank_screen_parts(nq, nb, min_parts, byref(parts)))
        parts = int(parts.value)
        q16 = queries32.to(torch.bfloat16).contiguous()
        cand = self._scratch("screen", parts * nq * 8 * 12, dev)
        cs = cand[:parts * nq * 8 * 4].view(torch.float32).view(parts, nq, 8)
        ci = cand[parts * nq * 8 * 4:parts * nq * 8 * 12].view(torch.int64).view(parts, nq, 8)
        scores = torch.empty((nq, k), dtype=torch.float32, device=dev)
        ids = torch.empty((nq, k), dtype=torch.int64, device=dev)
        unsure = torch.empty(nq, dtype=torch.int32, device=dev)
        n_unsure = torch.empty(1, dtype=torch.int32, device=dev)
        # |<q,b> - <bf16(q),bf16(b)>| <= (2 * 2^-9 + 2^-18) |q||b|; 1.05 covers the fp32 accumulation of both sides
        eps_scale = 1.05 * (2.0 ** -8 + 2.0 ** -18) * float(max_row_norm)
        with torch.cuda.device(dev):
            check(lib().bsc_bank_screen(q16.data_ptr(), nq, bank16.data_ptr(), nb, d, base_id, min_parts, cs.data_ptr(),
                                        ci.data_ptr(), stream_ptr(dev)))
            check(lib().bsc_bank_rescore(queries32.data_ptr(), nq, bank32.data_ptr(), nb, d, base_id, cs.data_ptr(),
                                         ci.data_ptr(), parts, k, eps_scale, scores.data_ptr(), ids.data_ptr(),
                                         unsure.data_ptr(), n_unsure.data_ptr(), stream_ptr(dev)))
        return scores, ids, unsure, n_unsure

    def bank_search_exact_finish(self, launched, queries32, bank32, k: int, base_id: int):
        """Second half of the exact search: one host read of the unproven count, exhaustive re-search of those."""
        scores, ids, unsure, n_unsure = launched
        n_bad = int(n_unsure.item())
        if n_bad:
            rows = torch.nonzero(unsure, as_tuple=False).flatten()
            s2, i2 = self.bank_search(queries32[rows].contiguous(), bank32, k, base_id, MODE_F32)
            scores[rows] = s2
            ids[rows] = i2
        return scores, ids, n_bad

    def bank_search_exact(self, queries32: torch.Tensor, bank32: torch.Tensor, bank16: torch.Tensor, k: int,
                          base_id: int, max_row_norm: float, min_parts: int = 16):
        """Exact float32 top-k through the bf16 screening pass + float32 re-scoring; queries whose result the bound
        cannot prove are re-run with the exhaustive float32 kernel.  Returns (scores, ids, n_rechecked)."""
        launched = self.bank_search_exact_launch(queries32, bank32, bank16, k, base_id, max_row_norm, min_parts)
        if launched is None:
            s, i = self.bank_search(queries32, bank32, k, base_id, MODE_F32)
            return s, i, queries32.shape[0]
        return self.bank_search_exact_finish(launched, queries32, bank32, k, base_id)

    def topk_merge(self, cand_scores: torch.Tensor, cand_ids: torch.Tensor):
        """[parts, nq, k] candidate lists -> [nq, k] (score desc, id asc)."""
        parts, nq, k = cand_scores.shape
        dev = cand_scores.device
        _lib.require_device(dev)
        cand_scores = cand_scores.contiguous()
        cand_ids = cand_ids.contiguous()
        scores = torch.empty((nq, k), dtype=torch.float32, device=dev)
        ids = torch.empty((nq, k), dtype=torch.int64, device=dev)
        with torch.cuda.device(dev):
            check(lib().bsc_topk_merge(cand_scores.data_ptr(), cand_ids.data_ptr(), parts, nq, k, scores.data_ptr(),
                                       ids.data_ptr(), stream_ptr(dev)))
        return scores, ids
