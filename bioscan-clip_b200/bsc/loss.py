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

from ._lib import MODE_F32, MODE_TC16, MODE_TC16X3, TC_MODES
from .ops import Problem, requested_bwd_mode

TOO_FEW = "Too less element for calculating the contrastive loss."  # loss_func.py:35,142

# Phase marks for tools/dist_phases.py: when a list is installed here, every phase boundary of a loss evaluation
# records a CUDA event on the current stream (instrumentation only; None in normal use).
PHASE_MARKS = None


def _mark(name: str) -> None:
    if PHASE_MARKS is not None:
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        PHASE_MARKS.append((name, ev))


class LocalComm:
    """world_size == 1: no exchange; the scatter strategy's exchange buffer is ordinary device memory."""
    rank = 0
    world_size = 1
    peer_memory = True

    def all_gather_rows(self, t: torch.Tensor) -> torch.Tensor:
        return t

    def share_stats(self, record: torch.Tensor) -> torch.Tensor:
        return record[None]

    def exchange(self, nbytes: int, device):
        buf = torch.empty(max(nbytes, 16), dtype=torch.uint8, device=device)
        return buf, [buf.data_ptr()]

    def exchange_barrier(self, device) -> None:
        return None


class TorchDistComm:
    """torch.distributed plumbing (NCCL on the GPU box, gloo in the CPU tests).  No peer-mapped memory: the backward
    uses the strategy that needs no exchange (weight blocks per ordered pair)."""
    peer_memory = False

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

    def share_stats(self, record: torch.Tensor) -> torch.Tensor:
        """Every rank's statistics record, [world, len] in rank order (bsc_infonce_loss adds them in that order: the
        result is the same on every rank and from run to run, which an all_reduce does not promise)."""
        out = torch.empty(self.world_size * record.shape[0], dtype=record.dtype, device=record.device)
        self.dist.all_gather_into_tensor(out, record.contiguous(), group=self.group)
        return out.view(self.world_size, record.shape[0])


class SymmMemComm(TorchDistComm):
    """TorchDistComm over NVLink peer memory (torch symmetric memory, one node):

    * the TC16 gather is ONE kernel: ``bsc_gather_push`` prepares the fp16 operands and stores each local row -- fp16
      values, scale, label -- into every rank's buffer, then a symmetric-memory barrier on the stream.  Two buffers
      alternate by forward: a rank starts pushing forward k+2 only after the barrier of forward k+1, which every rank
      passes after its forward-k kernels.  A problem whose backward is still outstanding when its buffer is about to
      be reused (two forwards before one backward, evaluation forwards in between) is moved to private copies first.
    * the scatter backward's exchange buffers (``bsc_infonce_bwd_partials`` stores partial gradients straight into the
      owners' buffers) live there too, two alternating by backward, with one barrier between the two phases.

    ``strict=False`` ('auto'): if the set-up fails on ANY rank all ranks agree (one all_reduce at set-up time) to
    continue on plain NCCL collectives; ``strict=True`` ('on') raises instead."""
    MAX_SHAPES = 2  # symmetric allocations are never returned to a pool: keep the cache of shapes small

    def __init__(self, rank: int, world_size: int, group=None, strict: bool = False):
        super().__init__(rank, world_size, group)
        self.strict = strict
        self.gather_path = "symm_mem"
        self._side = None
        self.disabled = False
        self._gather = {}     # (n_global, d, device) -> state
        self._exchange = {}   # (nbytes, device) -> state

    @property
    def peer_memory(self) -> bool:
        return not self.disabled and self.world_size <= 16

    # -- set-up with a collective verdict ---------------------------------------------------------
    def _agree(self, ok: bool, device) -> bool:
        flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=device)
        self.dist.all_reduce(flag, op=self.dist.ReduceOp.MIN, group=self.group)
        return bool(int(flag.item()))

    def _symm_buffers(self, nbytes: int, device, count: int):
        """``count`` symmetric uint8 buffers of nbytes, rendezvoused; None when any rank failed (non-strict)."""
        import warnings

        import torch.distributed._symmetric_memory as symm

        group = self.group if self.group is not None else self.dist.group.WORLD
        err = None
        bufs = []
        try:
            bufs = [symm.empty(int(nbytes), dtype=torch.uint8, device=device) for _ in range(count)]
        except Exception as e:  # noqa: BLE001 -- reported below, decided collectively
            err = e
        if not self._agree(err is None, device):
            if self.strict:
                raise err if err is not None else RuntimeError("symmetric memory allocation failed on a peer")
            warnings.warn(f"bsc: NVLink peer memory unavailable ({err!r} on this or another rank); using NCCL collectives")
            self.disabled = True
            return None
        handles, peers = [], []
        try:
            for t in bufs:
                h = symm.rendezvous(t, group)
                ptrs = list(h.buffer_ptrs)
                if len(ptrs) != self.world_size or h.rank != self.rank:
                    raise RuntimeError("symmetric memory group does not match the ClipLoss rank / world_size")
                handles.append(h)
                peers.append(ptrs)
        except Exception as e:  # noqa: BLE001
            err = e
        if not self._agree(err is None, device):
            if self.strict:
                raise err if err is not None else RuntimeError("symmetric memory rendezvous failed on a peer")
            warnings.warn(f"bsc: NVLink peer mapping unavailable ({err!r} on this or another rank); using NCCL collectives")
            self.disabled = True
            return None
        return bufs, handles, peers

    def _cache_put(self, cache: dict, key, value) -> None:
        while len(cache) >= self.MAX_SHAPES:
            cache.pop(next(iter(cache)))
        cache[key] = value

    # -- forward gather -----------------------------------------------------------------------------
    def gather_tc16(self, feats, labels):
        """The forward's exchange over peer memory.  Returns None when the inputs do not qualify (strided elements,
        unsupported dtype) or peer memory is unavailable -- the caller takes the NCCL path -- else a dict:
          xh[3], inv[3], labels   views of this rank's symmetric buffer (the gathered global batch)
          bind(problem)           registers the problem that will read them in its backward
          exchange                argument of CudaOps.infonce_fwd: where the statistics records go, and -- in the
                                  overlapped mode -- the arrival counters the sweep waits on
          records()               call after the forward sweep: barrier, then the [world, len] records of all ranks

        Overlapped mode (row blocks that are whole column tiles; BSC_GATHER_OVERLAP=0 turns it off): the labels go first
        (tiny, one barrier), the rows are prepared in this rank's own buffer, then a SIDE stream lets the copy engines
        carry them to one destination at a time while the main stream already runs the label statistics and the
        similarity sweep, which waits per source rank on arrival flags and starts with this rank's own columns.  Every GPU has to receive (W-1)/W of the batch (0.22 ms of NVLink ingress at
        8 x 4096 rows); this hides it behind the sweep instead of adding it to the step."""
        import ctypes
        import os
        import weakref

        from . import _lib

        if self.disabled:
            return None
        present = [f for f in feats if f is not None]
        f0 = present[0]
        if any(f.dtype != f0.dtype or f.stride(1) != 1 for f in present) or f0.shape[1] % 8 != 0:
            return None
        if f0.dtype not in (torch.float32, torch.bfloat16, torch.float16) or self.world_size > 16:
            return None
        n_local, d = f0.shape
        n_global = n_local * self.world_size
        dev = f0.device
        key = (n_global, d, dev)
        st = self._gather.get(key)
        L = _lib.lib()
        if st is None:
            offx, offi = (ctypes.c_size_t * 3)(), (ctypes.c_size_t * 3)()
            offl, offs, offf, total = (ctypes.c_size_t(0) for _ in range(4))
            _lib.check(L.bsc_gather_layout(n_global, d, self.world_size, offx, offi, ctypes.byref(offl),
                                           ctypes.byref(offs), ctypes.byref(offf), ctypes.byref(total)))
            made = self._symm_buffers(int(total.value), dev, 2)
            if made is None:
                return None
            bufs, handles, peers = made
            for t in bufs:
                t[int(offf.value):].zero_()  # arrival counters start at 0 and only ever grow
            torch.cuda.synchronize(dev)
            self.dist.barrier(group=self.group)  # nobody pushes before every counter is zeroed
            # NVSwitch multicast address of each buffer (0 / absent when the fabric has no NVLS): one multimem.st
            # reaches every rank.  All ranks see the same capability, so the choice needs no agreement.
            mcast = [int(getattr(h, "multicast_ptr", 0) or 0) for h in handles]
            if os.environ.get("BSC_GATHER_MULTICAST") == "0":
                mcast = [0 for _ in handles]
            st = {"bufs": bufs, "handles": handles, "mcast": mcast, "ptrs": peers,
                  "peers": [(ctypes.c_void_p * self.world_size)(*pp) for pp in peers],
                  "offx": list(offx), "offi": list(offi), "offl": int(offl.value), "offs": int(offs.value),
                  "offf": int(offf.value), "step": 0, "pending": [None, None], "pushes": [0, 0],
                  "stats_len": int(L.bsc_infonce_stats_len(n_global))}
            self._cache_put(self._gather, key, st)
        b = st["step"] & 1
        st["step"] += 1
        overlap = n_local % 256 == 0 and os.environ.get("BSC_GATHER_OVERLAP") != "0"
        self.gather_path = "symm_mem" + ("+overlapped" if overlap else "") + ("+multicast" if st["mcast"][b] else "")
        # the OTHER buffer is overwritten by the peers' next forward, which they may start as soon as this rank has
        # arrived at the barrier below: a problem still waiting for its backward there moves to private memory now
        for slot in (b ^ 1, b):
            ref = st["pending"][slot]
            prob = ref() if ref is not None else None
            if prob is not None and not getattr(prob, "done", False):
                prob.rebind([None if t is None else t.clone() for t in prob.xh],
                            [None if t is None else t.clone() for t in prob.inv], prob.all_labels.clone())
            st["pending"][slot] = None
        labels = labels.detach().to(torch.int64).contiguous()
        x = (ctypes.c_void_p * 3)(*[None if f is None else f.data_ptr() for f in feats])
        ldx = (ctypes.c_int64 * 3)(*[0 if f is None else f.stride(0) for f in feats])
        handle, mc = st["handles"][b], st["mcast"][b] or None
        exchange = {"stats_peers": [pp + st["offs"] for pp in st["ptrs"][b]],
                    "stats_multicast": (st["mcast"][b] + st["offs"]) if st["mcast"][b] else 0}
        side_done = None
        with torch.cuda.device(dev):
            main = torch.cuda.current_stream(dev)
            if overlap:
                _lib.check(L.bsc_gather_push_labels(labels.data_ptr(), n_local, self.rank * n_local, n_global, d,
                                                    st["peers"][b], mc, self.world_size, main.cuda_stream))
                handle.barrier(channel=0)
                _mark("labels_barrier")
                # own rows into the own buffer + own arrival flag, on the MAIN stream: done before the sweep starts
                st["pushes"][b] += 1
                _lib.check(L.bsc_gather_push(x, _lib.dtype_code(f0.dtype), ldx, None, n_local, self.rank * n_local,
                                             n_global, d, st["peers"][b], None, self.world_size, self.rank, 1,
                                             st["pushes"][b], main.cuda_stream))
                ready = torch.cuda.Event()
                ready.record(main)
                if self._side is None:
                    self._side = torch.cuda.Stream(dev)
                mask = sum(1 << m for m, f in enumerate(feats) if f is not None)
                with torch.cuda.stream(self._side):
                    # copy engines only (no kernel could be scheduled beside the persistent sweep): rows, inverse
                    # norms and the arrival flag to rank+1, rank+2, ...
                    self._side.wait_event(ready)
                    _lib.check(L.bsc_gather_ring(st["peers"][b], self.world_size, self.rank, n_local,
                                                 self.rank * n_local, n_global, d, mask, self._side.cuda_stream))
                    side_done = torch.cuda.Event()
                    side_done.record(self._side)
                exchange["arrive"] = st["ptrs"][b][self.rank] + st["offf"]
                exchange["arrive_target"] = st["pushes"][b]
                _mark("own_rows")
            else:
                _lib.check(L.bsc_gather_push(x, _lib.dtype_code(f0.dtype), ldx, labels.data_ptr(), n_local,
                                             self.rank * n_local, n_global, d, st["peers"][b], mc, self.world_size,
                                             self.rank, 0, 0, main.cuda_stream))
                _mark("gather_push")
                handle.barrier(channel=0)
                _mark("gather_barrier")
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

        def bind(prob, st=st, b=b):
            st["pending"][b] = weakref.ref(prob)

        def records(st=st, b=b, handle=handle, side_done=side_done, keep=feats):
            """After the sweep: every rank's record has been stored into every rank's area -> barrier -> read."""
            with torch.cuda.device(dev):
                if side_done is not None:
                    torch.cuda.current_stream(dev).wait_event(side_done)  # long done: the sweep saw every counter
                handle.barrier(channel=1)
            n = st["stats_len"]
            return st["bufs"][b][st["offs"]:st["offs"] + self.world_size * n * 4].view(torch.float32).view(
                self.world_size, n)

        return {"xh": xh, "inv": inv, "labels": all_labels, "bind": bind, "exchange": exchange, "records": records}

    # -- backward exchange ----------------------------------------------------------------------------
    def prepare_exchange(self, nbytes: int, device) -> bool:
        """Allocate and map the scatter backward's exchange buffers (collective).  False: peer memory is unavailable
        and every rank must plan again without it."""
        if self.disabled:
            return False
        key = (int(nbytes), device)
        if key in self._exchange:
            return True
        made = self._symm_buffers(int(nbytes), device, 2)
        if made is None:
            return False
        bufs, handles, peers = made
        self._cache_put(self._exchange, key, {"bufs": bufs, "handles": handles, "peers": peers, "step": 0})
        return True

    def exchange(self, nbytes: int, device):
        st = self._exchange.get((int(nbytes), device))
        if st is None:
            raise RuntimeError("scatter backward without prepared exchange buffers (prepare_exchange runs in the forward)")
        b = st["step"] & 1
        st["step"] += 1
        self._ex_handle = st["handles"][b]
        return st["bufs"][b], st["peers"][b]

    def exchange_barrier(self, device) -> None:
        with torch.cuda.device(device):
            self._ex_handle.barrier(channel=1)


X3_MIN_ROWS = 2048  # 'auto' with float32 inputs: batches from this size on take the split-operand tensor-core mode


def resolve_mode(precision: str, feats: Sequence[torch.Tensor], n_global: Optional[int] = None) -> int:
    """'fp32'   SIMT float32 tiles (BSC_MODE_F32): the small-batch checker, any width;
    'fp32x3' float32-grade arithmetic on the tensor cores (BSC_MODE_TC16X3: hi + lo fp16 parts, three MMA terms);
    'tc16'   the 16-bit throughput mode (BSC_MODE_TC16);
    'auto'   the reference computes in the input dtype: float32 inputs take a float32-grade mode ('fp32x3' from
             X3_MIN_ROWS rows of the global batch on, 'fp32' below), 16-bit inputs 'tc16'; widths the tensor-core
             kernels do not tile take 'fp32'."""
    d = feats[0].shape[1]
    if precision == "fp32":
        return MODE_F32
    if precision in ("tc16", "bf16", "fp16", "fp32x3"):
        if d % 128 != 0:
            from ._lib import BscError

            raise BscError(f"the tensor-core modes need an embedding dimension that is a multiple of 128 (got {d}); "
                           "use precision='fp32' (BSC_MODE_F32) or 'auto'")
        return MODE_TC16X3 if precision == "fp32x3" else MODE_TC16
    if precision != "auto":
        raise ValueError(f"precision must be 'auto', 'fp32', 'fp32x3' or 'tc16' (got {precision!r})")
    if d % 128 != 0:
        return MODE_F32
    if all(f.dtype == torch.float32 for f in feats):
        rows = n_global if n_global is not None else feats[0].shape[0]
        return MODE_TC16X3 if rows >= X3_MIN_ROWS else MODE_F32
    return MODE_TC16


def infonce_forward(ops, comm, feats: List[Optional[torch.Tensor]], labels: torch.Tensor, scale: float, mode: int,
                    want_grad: bool = True, bwd_request: int = -1):
    """Forward of the fused loss.  Returns (loss, saved-for-backward)."""
    present = [f for f in feats if f is not None]
    n_local, d = present[0].shape
    for f in present:
        if f.shape != (n_local, d):
            raise ValueError("all modalities must have the same [batch, dim] shape")
        if f.dtype != present[0].dtype:
            raise ValueError("all modalities must share one dtype (the reference computes in one dtype, float32)")
    if labels.shape[0] != n_local:
        raise ValueError("label batch does not match the feature batch")
    inv = None
    all_labels = None
    bind = None
    fused = None
    _mark("start")
    if mode == MODE_F32:
        xh = [None if f is None else comm.all_gather_rows(ops.normalize(f.detach(), torch.float32)) for f in feats]
    elif mode == MODE_TC16X3:
        # float32-grade tensor-core mode: hi | lo fp16 parts of every row and the float32 inverse norms, gathered with
        # plain collectives (the fused NVLink gather carries BSC_MODE_TC16 rows only)
        xh, inv = [], []
        for f in feats:
            if f is None:
                xh.append(None)
                inv.append(None)
                continue
            hl, iv = ops.rows_to_f16x3(f.detach())
            xh.append(comm.all_gather_rows(hl))
            inv.append(comm.all_gather_rows(iv))
    else:
        # tensor-core mode: the MMA operands are the caller's values times an exact per-row power of two, held in
        # float16 (exact for bf16 / fp16 inputs; float32 inputs are rounded once to 11 bits).  Normalised rows are
        # never re-rounded: the inverse norms enter as float32 scales inside the kernels.
        fused = comm.gather_tc16([None if f is None else f.detach() for f in feats], labels) \
            if hasattr(comm, "gather_tc16") else None
        if fused is not None:
            xh, inv, all_labels = fused["xh"], fused["inv"], fused["labels"]
            bind = fused.get("bind")
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
    _mark("gather")
    r, rep, nxt, rmax = ops.label_stats(all_labels)
    _mark("label_stats")
    prob = Problem(xh=xh, n_global=n_global, n_local=n_local, row0=comm.rank * n_local, d=d, mode=mode,
                   scale=scale, r=r, rep=rep, nxt=nxt, rmax=rmax, inv=inv, all_labels=all_labels,
                   world=comm.world_size, rank=comm.rank)
    ops.plan(prob, bwd_request, peer_memory=bool(getattr(comm, "peer_memory", False)))
    if want_grad and prob.exchange_bytes and comm.world_size > 1:
        if not comm.prepare_exchange(prob.exchange_bytes, present[0].device):  # collective verdict
            ops.plan(prob, bwd_request, peer_memory=False)
    if bind is not None and want_grad:
        bind(prob)
    if fused is not None and fused.get("records") is not None:
        # the sweep stores this rank's statistics record straight into every rank's record area (peer memory)
        ops.infonce_fwd(prob, fused.get("exchange"))
        _mark("fwd_sweep")
        records = fused["records"]()
    else:
        record = ops.infonce_fwd(prob)
        _mark("fwd_sweep")
        records = comm.share_stats(record)
    _mark("stats_exchange")
    loss, stats, aux = ops.infonce_loss(prob, records)
    _mark("loss")
    return loss, (prob, aux)


def infonce_backward(ops, comm, saved, feats, grad_out: torch.Tensor, grad_gain: float = 1.0):
    from ._lib import BWD_SCATTER

    prob, aux = saved
    if grad_gain != 1.0:
        grad_out = grad_out * grad_gain
    xs = [None if f is None else f.detach() for f in feats]
    dev = next(x for x in xs if x is not None).device
    own, ptrs = (None, None)
    if prob.mode in TC_MODES and prob.bwd_mode == BWD_SCATTER:
        own, ptrs = comm.exchange(prob.exchange_bytes, dev)
    _mark("bwd_start")
    ops.infonce_bwd_partials(prob, aux, ptrs)
    _mark("bwd_partials")
    if own is not None:
        comm.exchange_barrier(dev)  # every rank's partial sums have landed before anyone adds them up
        _mark("exchange_barrier")
    grads = ops.infonce_bwd_finish(prob, aux, xs, grad_out, own)
    _mark("bwd_finish")
    prob.done = True
    prob.bwd_workspace = None
    return grads


class _InfoNCEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, ops, comm, scale, mode, grad_gain, bwd_request, labels, image, dna, text):
        feats = [image, dna, text]
        want_grad = any(f is not None and f.requires_grad for f in feats)
        loss, saved = infonce_forward(ops, comm, feats, labels, scale, mode, want_grad, bwd_request)
        ctx.ops = ops
        ctx.comm = comm
        ctx.saved_state = saved
        ctx.save_for_backward(image, dna, text)
        ctx.grad_gain = grad_gain
        return loss

    @staticmethod
    def backward(ctx, grad_out):
        feats = list(ctx.saved_tensors)
        if ctx.saved_state is None:
            raise RuntimeError("bsc loss: backward called twice on the same graph; evaluate the loss again instead "
                               "(the operands of the first pass are released after its backward)")
        grads = infonce_backward(ctx.ops, ctx.comm, ctx.saved_state, feats, grad_out, ctx.grad_gain)
        ctx.saved_state = None
        out = []
        for f, g in zip(feats, grads):
            out.append(g if (f is not None and f.requires_grad) else None)
        return (None, None, None, None, None, None, None, *out)


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


def _plain_scale(logit_scale) -> float:
    """The temperature is a constant in the reference (scripts/train_cl.py:190; the learnable one is commented out,
    simple_clip.py:54-79): a tensor that wants a gradient would silently lose it here, so it is refused."""
    if isinstance(logit_scale, torch.Tensor):
        if logit_scale.requires_grad:
            raise NotImplementedError("bsc losses take a constant logit_scale (no gradient flows to it)")
        return float(logit_scale.detach())
    return float(logit_scale)


def _default_ops():
    from .ops import CudaOps

    return CudaOps()


class ContrastiveLoss(nn.Module):
    """Single-process multi-modal InfoNCE with multi-positive labels (loss_func.py:23-54)."""

    def __init__(self, criterion=None, logit_scale=1 / 0.07, precision: str = "auto", backward: str = "auto"):
        # precision: 'auto' | 'fp32' | 'fp32x3' | 'tc16' (resolve_mode)
        super().__init__()
        _check_criterion(criterion)
        self.criterion = criterion
        self.logit_scale = _plain_scale(logit_scale)
        self.precision = precision
        self.backward_mode = backward  # auto | scatter | blocks | recompute (BSC_BWD_MODE overrides)
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
        return _InfoNCEFn.apply(self._ops, LocalComm(), float(self.logit_scale), mode, 1.0,
                                requested_bwd_mode(self.backward_mode), label, *feats)


class ClipLoss(nn.Module):
    """Distributed twin (loss_func.py:94-165): every rank returns the loss of the GLOBAL batch."""

    def __init__(self, local_loss=False, gather_with_grad=False, cache_labels=False, rank=0, world_size=1,
                 use_horovod=False, criterion=None, precision: str = "auto", group=None, fused_gather: str = "auto",
                 backward: str = "auto"):
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
        self.backward_mode = backward
        self._ops = None
        self._comm = None

    def _make_comm(self, device):
        """NCCL plumbing, with NVLink peer memory (SymmMemComm: one-kernel gather, scatter backward) on top when the
        ranks share a node.  'auto' continues on plain NCCL collectives when the set-up fails on any rank (decided
        collectively); 'on' raises; 'off' never tries."""
        import os

        want = self.fused_gather
        if os.environ.get("BSC_FUSED_GATHER") == "0":
            want = "off"
        if want == "off" or device.type != "cuda":
            return TorchDistComm(self.rank, self.world_size, self.group)
        return SymmMemComm(self.rank, self.world_size, self.group, strict=(want == "on"))

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
        mode = resolve_mode(self.precision, present, present[0].shape[0] * max(int(self.world_size), 1))
        gain = float(self.world_size) if (self.gather_with_grad and self.world_size > 1) else 1.0
        total = _InfoNCEFn.apply(self._ops, comm, _plain_scale(logit_scale), mode, gain,
                                 requested_bwd_mode(self.backward_mode), labels, *feats)
        return {"contrastive_loss": total} if output_dict else total
