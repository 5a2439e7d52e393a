"""Drop-in loss modules for bioscanclip/model/loss_func.py.

* ``ContrastiveLoss(criterion, logit_scale)``  -- loss_func.py:23-54, built at scripts/train_cl.py:190 and
  called as ``criterion(image_output, dna_output, language_output, label)`` (epoch/train_epoch.py:37).
* ``ClipLoss(local_loss, gather_with_grad, cache_labels, rank, world_size, use_horovod, criterion)``
  -- loss_func.py:94-165; ``forward(image, dna, text, labels, logit_scale, output_dict=False)``.

Same signatures, same ValueError text, a 0-dim float32 tensor on the autograd graph.  The arithmetic runs in
libbsc_hot.so (CUDA, sm_100a); nothing here falls back to PyTorch math.

Row partition for world_size > 1: rank r owns rows [r*n, (r+1)*n) of every modality.
Exchange: one all_gather of the normalised rows per modality plus the labels, then one all_reduce of the
6*N+4 statistics floats.  The backward needs no collective under the reference's default
``gather_with_grad=False`` (only the local slot carries gradient, loss_func.py:86-88); with
``gather_with_grad=True`` every rank's identical loss back-propagates into every rank's rows, i.e. the
gradient is world_size times larger, which is what the reference yields.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch
import torch.nn as nn

from ._lib import MODE_F32, MODE_TC16
from .ops import Problem

TOO_FEW = "Too less element for calculating the contrastive loss."  # loss_func.py:35,142


class LocalComm:
    """world_size == 1: no exchange."""
    rank = 0
    world_size = 1

    def all_gather_rows(self, t: torch.Tensor) -> torch.Tensor:
        return t

    def all_reduce_sum(self, t: torch.Tensor) -> torch.Tensor:
        return t


class TorchDistComm:
    """torch.distributed plumbing (NCCL on the GPU box, gloo in the CPU tests)."""

    def __init__(self, rank: int, world_size: int, group=None):
        import torch.distributed as dist

        self.dist = dist
        self.rank = rank
        self.world_size = world_size
        self.group = group

    def all_gather_rows(self, t: torch.Tensor) -> torch.Tensor:
        out = torch.empty((self.world_size * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
        self.dist.all_gather_into_tensor(out, t.contiguous(), group=self.group)
        return out

    def all_reduce_sum(self, t: torch.Tensor) -> torch.Tensor:
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM, group=self.group)
        return t


class SymmMemComm(TorchDistComm):
    """TorchDistComm whose TC16 gather is ONE kernel over NVLink peer memory instead of seven NCCL calls.

    Every rank owns two buffers of ``bsc_gather_layout`` bytes in torch symmetric memory (mapped into all ranks of
    the node).  ``bsc_gather_push`` prepares the fp16 operands and stores each local row -- fp16 values, scale,
    label -- into every rank's buffer; a symmetric-memory barrier on the stream follows.  The two buffers alternate
    by step: a rank only starts pushing step k+2 after the barrier of step k+1, which every rank passes after its
    step-k kernels (same stream), so no rank still reads the buffer being overwritten.
    """

    def __init__(self, rank: int, world_size: int, group=None):
        super().__init__(rank, world_size, group)
        self._state = {}

    def _setup(self, n_global: int, d: int, device):
        import ctypes

        import torch.distributed._symmetric_memory as symm

        from . import _lib

        offx, offi = (ctypes.c_size_t * 3)(), (ctypes.c_size_t * 3)()
        offl, total = ctypes.c_size_t(0), ctypes.c_size_t(0)
        _lib.check(_lib.lib().bsc_gather_layout(n_global, d, offx, offi, ctypes.byref(offl), ctypes.byref(total)))
        group = self.group if self.group is not None else self.dist.group.WORLD
        bufs, handles, peers = [], [], []
        for _ in range(2):
            t = symm.empty(int(total.value), dtype=torch.uint8, device=device)
            h = symm.rendezvous(t, group)
            ptrs = list(h.buffer_ptrs)
            if len(ptrs) != self.world_size or h.rank != self.rank:
                raise RuntimeError("symmetric memory group does not match the ClipLoss rank / world_size")
            bufs.append(t)
            handles.append(h)
            peers.append((ctypes.c_void_p * self.world_size)(*ptrs))
        return {"bufs": bufs, "handles": handles, "peers": peers, "offx": list(offx), "offi": list(offi),
                "offl": int(offl.value), "step": 0}

    def gather_tc16(self, feats, labels):
        """-> (xh[3], inv[3], all_labels) as views of this rank's symmetric buffer, or None when the inputs do not
        qualify (mixed dtypes, strided rows) and the caller must take the NCCL path."""
        import ctypes

        from . import _lib

        present = [f for f in feats if f is not None]
        f0 = present[0]
        if any(f.dtype != f0.dtype or f.stride(1) != 1 for f in present) or f0.shape[1] % 8 != 0:
            return None
        if f0.dtype not in (torch.float32, torch.bfloat16, torch.float16) or self.world_size > 16:
            return None
        n_local, d = f0.shape
        n_global = n_local * self.world_size
        key = (n_global, d, f0.device)
        st = self._state.get(key)
        if st is None:
            st = self._state[key] = self._setup(n_global, d, f0.device)
        b = st["step"] & 1
        st["step"] += 1
        labels = labels.detach().to(torch.int64).contiguous()
        x = (ctypes.c_void_p * 3)(*[None if f is None else f.data_ptr() for f in feats])
        ldx = (ctypes.c_int64 * 3)(*[0 if f is None else f.stride(0) for f in feats])
        with torch.cuda.device(f0.device):
            _lib.check(_lib.lib().bsc_gather_push(x, _lib.dtype_code(f0.dtype), ldx, labels.data_ptr(), n_local,
                                                  self.rank * n_local, n_global, d, st["peers"][b], self.world_size,
                                                  _lib.stream_ptr(f0.device)))
            st["handles"][b].barrier(channel=0)
        buf = st["bufs"][b]
        xh, inv = [], []
        for m, f in enumerate(feats):
            if f is None:
                xh.append(None)
                inv.append(None)
                continue
            xh.append(buf[st["offx"][m]:st["offx"][m] + n_global * d * 2].view(torch.float16).view(n_global, d))
            inv.append(buf[st["offi"][m]:st["offi"][m] + n_global * 4].view(torch.float32))
        all_labels = buf[st["offl"]:st["offl"] + n_global * 8].view(torch.int64)
        return xh, inv, all_labels


def resolve_mode(precision: str, feats: Sequence[torch.Tensor]) -> int:
    if precision == "fp32":
        return MODE_F32
    if precision in ("tc16", "bf16", "fp16"):
        return MODE_TC16
    if precision != "auto":
        raise ValueError(f"precision must be 'auto', 'fp32' or 'tc16' (got {precision!r})")
    # auto: the reference computes in the input dtype -> fp32 inputs take the exact path
    return MODE_F32 if all(f.dtype == torch.float32 for f in feats) else MODE_TC16


def infonce_forward(ops, comm, feats: List[Optional[torch.Tensor]], labels: torch.Tensor, scale: float, mode: int):
    """Forward of the fused loss.  Returns (loss, saved-for-backward)."""
    present = [f for f in feats if f is not None]
    n_local, d = present[0].shape
    for f in present:
        if f.shape != (n_local, d):
            raise ValueError("all modalities must have the same [batch, dim] shape")
    if labels.shape[0] != n_local:
        raise ValueError("label batch does not match the feature batch")
    inv = None
    all_labels = None
    if mode == MODE_F32:
        xh = [None if f is None else comm.all_gather_rows(ops.normalize(f.detach(), torch.float32)) for f in feats]
    else:
        # tensor-core mode: the MMA operands are the caller's values times an exact per-row power of two, held in
        # float16 (exact for bf16 / fp16 inputs; float32 inputs are rounded once to 11 bits).  Normalised rows are
        # never re-rounded: the inverse norms enter as float32 scales inside the kernels.
        fused = comm.gather_tc16([None if f is None else f.detach() for f in feats], labels) \
            if hasattr(comm, "gather_tc16") else None
        if fused is not None:
            xh, inv, all_labels = fused
        else:
            xh, inv = [], []
            for f in feats:
                if f is None:
                    xh.append(None)
                    inv.append(None)
                    continue
                y16, iv = ops.rows_to_f16(f.detach())
                xh.append(comm.all_gather_rows(y16))
                inv.append(comm.all_gather_rows(iv))
    if all_labels is None:
        all_labels = comm.all_gather_rows(labels.detach().to(torch.int64).contiguous())
    n_global = all_labels.shape[0]
    r, rep, nxt, rmax = ops.label_stats(all_labels)
    prob = Problem(xh=xh, n_global=n_global, n_local=n_local, row0=comm.rank * n_local, d=d, mode=mode,
                   scale=scale, r=r, rep=rep, nxt=nxt, rmax=rmax, inv=inv)
    stats = ops.infonce_fwd(prob)
    stats = comm.all_reduce_sum(stats)
    loss = ops.infonce_loss(prob, stats)
    return loss, (prob, stats)


def infonce_backward(ops, saved, feats, grad_out: torch.Tensor, grad_gain: float = 1.0):
    prob, stats = saved
    if grad_gain != 1.0:
        grad_out = grad_out * grad_gain
    return ops.infonce_bwd(prob, stats, [None if f is None else f.detach() for f in feats], grad_out)


class _InfoNCEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ops, comm, scale, mode, grad_gain, labels, image, dna, text):
        feats = [image, dna, text]
        loss, saved = infonce_forward(ops, comm, feats, labels, scale, mode)
        ctx.ops = ops
        ctx.saved_state = saved
        ctx.save_for_backward(image, dna, text)
        ctx.grad_gain = grad_gain
        return loss

    @staticmethod
    def backward(ctx, grad_out):
        feats = list(ctx.saved_tensors)
        grads = infonce_backward(ctx.ops, ctx.saved_state, feats, grad_out, ctx.grad_gain)
        ctx.saved_state = None
        out = []
        for f, g in zip(feats, grads):
            out.append(g if (f is not None and f.requires_grad) else None)
        return (None, None, None, None, None, None, *out)


def _check_criterion(criterion) -> None:
    """Only the configuration the reference constructs is implemented (scripts/train_cl.py:190)."""
    if criterion is None:
        return
    ok = isinstance(criterion, nn.CrossEntropyLoss) and criterion.reduction == "mean" and \
        criterion.label_smoothing == 0.0 and criterion.weight is None
    if not ok:
        raise NotImplementedError(
            "bsc losses implement nn.CrossEntropyLoss() with default arguments only (mean reduction, no "
            "label smoothing, no class weights), which is what the reference constructs")


def _default_ops():
    from .ops import CudaOps

    return CudaOps()


class ContrastiveLoss(nn.Module):
    """Single-process multi-modal InfoNCE with multi-positive labels (loss_func.py:23-54)."""

    def __init__(self, criterion=None, logit_scale=1 / 0.07, precision: str = "auto"):
        super().__init__()
        _check_criterion(criterion)
        self.criterion = criterion
        self.logit_scale = logit_scale
        self.precision = precision
        self._ops = None

    def forward(self, image_features, dna_features, text_features, label, logit_scale=1 / 0.07):
        # like the reference, the forward kwarg is ignored in favour of self.logit_scale (loss_func.py:29 vs :46)
        feats = [image_features, dna_features, text_features]
        present = [f for f in feats if f is not None]
        if len(present) < 2:
            raise ValueError(TOO_FEW)
        if self._ops is None:
            self._ops = _default_ops()
        mode = resolve_mode(self.precision, present)
        return _InfoNCEFn.apply(self._ops, LocalComm(), float(self.logit_scale), mode, 1.0, label, *feats)


class ClipLoss(nn.Module):
    """Distributed twin (loss_func.py:94-165): every rank returns the loss of the GLOBAL batch."""

    def __init__(self, local_loss=False, gather_with_grad=False, cache_labels=False, rank=0, world_size=1,
                 use_horovod=False, criterion=None, precision: str = "auto", group=None, fused_gather: str = "auto"):
        super().__init__()
        if use_horovod:
            raise NotImplementedError("horovod gathers are not implemented; use torch.distributed (NCCL)")
        _check_criterion(criterion)
        self.local_loss = local_loss
        self.gather_with_grad = gather_with_grad
        self.cache_labels = cache_labels
        self.rank = rank
        self.world_size = world_size
        self.use_horovod = use_horovod
        self.criterion = criterion
        self.precision = precision
        self.group = group
        if fused_gather not in ("auto", "on", "off"):
            raise ValueError("fused_gather must be 'auto', 'on' or 'off'")
        self.fused_gather = fused_gather
        self._ops = None
        self._comm = None

    def _make_comm(self, device):
        """NCCL plumbing, with the one-kernel NVLink gather (SymmMemComm) on top when the ranks share a node's
        symmetric memory.  'auto' falls back to plain NCCL collectives (still on the GPU) if the set-up fails."""
        import os
        import warnings

        want = self.fused_gather
        if os.environ.get("BSC_FUSED_GATHER") == "0":
            want = "off"
        if want == "off" or device.type != "cuda":
            return TorchDistComm(self.rank, self.world_size, self.group)
        comm = SymmMemComm(self.rank, self.world_size, self.group)
        if want == "on":
            return comm
        inner = comm.gather_tc16

        def guarded(feats, labels):
            try:
                return inner(feats, labels)
            except Exception as e:  # set-up or peer mapping failed: keep training on the NCCL collectives
                warnings.warn(f"bsc: fused NVLink gather unavailable ({type(e).__name__}: {e}); using NCCL all_gather")
                comm.gather_tc16 = lambda feats, labels: None
                return None

        comm.gather_tc16 = guarded
        return comm

    def forward(self, image_features, dna_features, text_features, labels, logit_scale, output_dict=False):
        feats = [image_features, dna_features, text_features]
        present = [f for f in feats if f is not None]
        if len(present) < 2:
            raise ValueError(TOO_FEW)
        if self._ops is None:
            self._ops = _default_ops()
        if self._comm is None:
            self._comm = self._make_comm(present[0].device) if self.world_size > 1 else LocalComm()
        comm = self._comm
        mode = resolve_mode(self.precision, present)
        gain = float(self.world_size) if (self.gather_with_grad and self.world_size > 1) else 1.0
        total = _InfoNCEFn.apply(self._ops, comm, float(logit_scale), mode, gain, labels, *feats)
        return {"contrastive_loss": total} if output_dict else total
