"""GPU parity: the CUDA loss (through the C ABI) vs the oracle and vs the committed reference fixtures."""
import glob
import os

import numpy as np
import pytest
import torch

from oracle import loss_oracle
from tests.synth import loss_inputs

pytestmark = pytest.mark.gpu
NAMES = ["image", "dna", "text"]
FP32_TOL = 1e-5   # north_star: loss and gradients within 1e-5 relative in fp32
TC16_TOL = 2e-3   # and 2e-3 in the 16-bit tensor-core mode


def _run(feats, labels, precision, scale=1 / 0.07):
    import bsc

    leaves = [None if f is None else f.detach().clone().cuda().requires_grad_(True) for f in feats]
    crit = bsc.ContrastiveLoss(torch.nn.CrossEntropyLoss(), scale, precision=precision)
    loss = crit(leaves[0], leaves[1], leaves[2], labels.cuda())
    loss.backward()
    torch.cuda.synchronize()
    return loss.detach().cpu(), [None if l is None else l.grad.detach().float().cpu() for l in leaves]


LOSS_FLOOR = 5e-7  # half an ulp of a float32 logit near s = 14.3: the reference's own logits carry this much


def _compare(loss, grads, ref_loss, ref_grads, tol, noise_loss=0.0, noise_grads=None):
    """|ours - ref| <= tol * |ref|  (+ the float32 rounding noise of the reference arithmetic itself, measured
    as the distance between a float32 and a float64 evaluation of the reference on the same inputs)."""
    bound = tol * abs(float(ref_loss)) + 2.0 * noise_loss + LOSS_FLOOR
    assert abs(float(loss) - float(ref_loss)) <= bound, (float(loss), float(ref_loss), bound)
    for k, (g, gr) in enumerate(zip(grads, ref_grads)):
        if gr is None:
            assert g is None
            continue
        err = (g.double() - gr.double()).abs().max().item()
        noise = 0.0 if noise_grads is None else noise_grads[k]
        bound = tol * gr.abs().max().item() + 2.0 * noise + 1e-12
        assert err <= bound, (k, err, gr.abs().max().item(), bound)


def _fp32_noise(feats, labels, ref_loss64, ref_grads64, scale=1 / 0.07):
    """Distance between the float32 reference arithmetic (oracle port, autograd) and the float64 closed form."""
    l32, g32 = loss_oracle.contrastive_loss_and_grads(feats, labels, scale)
    nl = abs(float(l32) - float(ref_loss64))
    ng = [0.0 if g is None else (g.double() - r.double()).abs().max().item() for g, r in zip(g32, ref_grads64)]
    return nl, ng


@pytest.mark.parametrize("path", sorted(glob.glob(os.path.join(os.path.dirname(__file__), "golden", "loss_*.npz"))))
def test_fp32_mode_matches_reference_fixtures(path):
    z = np.load(path)
    feats = [torch.from_numpy(z[f"x_{n}"]) if f"x_{n}" in z else None for n in NAMES]
    labels = torch.from_numpy(z["labels"])
    scale = float(z["logit_scale"])
    loss, grads = _run(feats, labels, "fp32", scale)
    ref_grads = [torch.from_numpy(z[f"g_{n}"]) if f"g_{n}" in z else None for n in NAMES]
    # the fixture is the reference's own float32 output: its distance from the float64 closed form on the same
    # inputs is the rounding noise of the reference arithmetic
    l64, g64 = loss_oracle.closed_form_grads(feats, labels, scale)
    nl = abs(float(z["loss"]) - float(l64))
    ng = [0.0 if g is None else (g.double() - r.double()).abs().max().item() for g, r in zip(ref_grads, g64)]
    _compare(loss, grads, float(z["loss"]), ref_grads, FP32_TOL, nl, ng)


@pytest.mark.parametrize("n,d,present,labels,normalised", [
    (256, 768, (True, True, False), "arange", True),    # BASELINE config 1
    (300, 200, (True, True, True), "multi", False),     # ragged N, d not a multiple of 16
    (129, 768, (True, False, True), "multi", True),
    (1024, 768, (True, True, True), "arange", True),
    (64, 32, (True, True, True), "same", False),        # every label equal: r_i = N
])
def test_fp32_mode_matches_oracle(n, d, present, labels, normalised):
    feats, lab = loss_inputs(n, d, present, seed=n, normalised=normalised, labels=labels)
    ref_loss, ref_grads = loss_oracle.closed_form_grads(feats, lab)  # fp64 evaluation of the reference maths
    nl, ng = _fp32_noise(feats, lab, ref_loss, ref_grads)
    loss, grads = _run(feats, lab, "fp32")
    _compare(loss, grads, ref_loss, ref_grads, FP32_TOL, nl, ng)


def _bf16_round_bound(ref):
    """One rounding of the result to bfloat16 (8-bit significand): at most 2^-8 of the value."""
    return ref.abs() * 2.0 ** -8


TC16_CASES = [
    (256, 768, (True, True, False), "arange", 0.5),
    (512, 768, (True, True, True), "multi", 0.5),
    (1000, 768, (True, True, True), "arange", 0.5),     # ragged N (not a tile multiple)
    (384, 256, (True, False, True), "multi", 0.0),      # flat softmax, d = 2 column groups
    (2048, 768, (True, True, True), "arange", 0.5),
    (640, 512, (True, True, True), "multi", 0.5),       # odd number of column tiles (pair sweep pads one), d = 2 x 256
    (130, 768, (False, True, True), "multi", 0.5),      # two row blocks, the second almost empty
]


@pytest.mark.parametrize("n,d,present,labels,shared", TC16_CASES)
@pytest.mark.parametrize("io", ["float32", "bfloat16", "float16"])
def test_tc16_mode_matches_oracle_on_16bit_inputs(n, d, present, labels, shared, io):
    """Tensor-core mode against the fp32 reference maths evaluated on the SAME 16-bit-representable inputs.
    io=float32: the rows hold bf16-representable values in float32 tensors, so the gradients come back in float32
    and the 2e-3 bound applies to the kernel arithmetic alone; io=bfloat16/float16: the gradients are additionally
    rounded once to the leaf dtype."""
    op = torch.float16 if io == "float16" else torch.bfloat16
    feats, lab = loss_inputs(n, d, present, seed=n + 1, labels=labels, shared=shared, dtype=op)
    ref_in = [None if f is None else f.float() for f in feats]
    ref_loss, ref_grads = loss_oracle.closed_form_grads(ref_in, lab)
    leaves = ref_in if io == "float32" else feats
    loss, grads = _run(leaves, lab, "tc16")
    assert abs(float(loss) - float(ref_loss)) <= TC16_TOL * abs(float(ref_loss)) + LOSS_FLOOR
    for g, gr in zip(grads, ref_grads):
        if gr is None:
            assert g is None
            continue
        err = (g.double() - gr.double()).abs()
        bound = TC16_TOL * gr.abs().max()
        if io != "float32":
            # float16 leaves: 11-bit significand, and gradients below 6.1e-5 land on the subnormal grid (spacing 2^-24)
            bound = bound + (_bf16_round_bound(gr) if io == "bfloat16" else gr.abs() * 2.0 ** -11 + 2.0 ** -25)
        assert (err <= bound + 1e-12).all(), (err.max().item(), gr.abs().max().item())


BWD_CODE = {"scatter": 4, "blocks": 3, "recompute": 0}


@pytest.mark.parametrize("n,d,present,labels,shared", TC16_CASES)
def test_tc16_backward_modes_match_oracle_and_each_other(n, d, present, labels, shared, monkeypatch):
    """Three backward strategies (include/bsc_hot.h BSC_BWD_*).  scatter (the default): the fp16 weight block of every
    UNORDERED pair is written once and serves both gradient GEMMs; blocks: one block per ORDERED pair, one GEMM each
    (the multi-GPU path without peer memory); recompute: the fused sweep that never stores anything N^2-sized.
    All three against the oracle, and against each other."""
    from bsc import _lib

    feats, lab = loss_inputs(n, d, present, seed=n + 1, labels=labels, shared=shared, dtype=torch.bfloat16)
    ref_in = [None if f is None else f.float() for f in feats]
    ref_loss, ref_grads = loss_oracle.closed_form_grads(ref_in, lab)
    results = {}
    for mode in ("scatter", "blocks", "recompute"):
        monkeypatch.setenv("BSC_BWD_MODE", mode)
        results[mode] = _run(ref_in, lab, "tc16")
        assert _lib.last_bwd_mode() == BWD_CODE[mode]
        _compare(results[mode][0], results[mode][1], ref_loss, ref_grads, TC16_TOL)
    for mode in ("scatter", "blocks"):
        assert float(results[mode][0]) == float(results["recompute"][0])  # the forward does not depend on the backward
        for ga, gb in zip(results[mode][1], results["recompute"][1]):
            if gb is None:
                assert ga is None
                continue
            assert (ga - gb).abs().max().item() <= 1.5e-3 * gb.abs().max().item()


@pytest.mark.parametrize("n,d,present", [(512, 128, (True, True, True)), (384, 384, (True, True, False)),
                                         (300, 1024, (True, True, True)), (256, 640, (False, True, True)),
                                         (200, 1152, (True, True, True))])
@pytest.mark.parametrize("mode", ["scatter", "blocks"])
def test_tc16_embedding_widths_other_than_768(n, d, present, mode, monkeypatch):
    """d = any multiple of 128: 384-wide items where d % 384 == 0, 256-wide otherwise (a last half-empty chunk when
    d % 256 == 128).  The recomputing sweep covers {256, 512, 768} only and must refuse the others loudly."""
    import bsc
    from bsc import _lib

    feats, lab = loss_inputs(n, d, present, seed=d, labels="multi", dtype=torch.bfloat16)
    ref_in = [None if f is None else f.float() for f in feats]
    ref_loss, ref_grads = loss_oracle.closed_form_grads(ref_in, lab)
    monkeypatch.setenv("BSC_BWD_MODE", mode)
    loss, grads = _run(ref_in, lab, "tc16")
    assert _lib.last_bwd_mode() == BWD_CODE[mode]
    _compare(loss, grads, ref_loss, ref_grads, TC16_TOL)
    monkeypatch.setenv("BSC_BWD_MODE", "recompute")
    with pytest.raises(bsc.BscError, match="256, 512, 768"):
        _run(ref_in, lab, "tc16")


def test_tc16_backward_falls_back_to_recomputing_sweep_above_the_memory_cap(monkeypatch):
    """The weight blocks are the one N x N-sized buffer of the design: when they do not fit the memory budget
    (free device memory, or BSC_BWD_MEM_GB) the backward recomputes instead."""
    from bsc import _lib

    feats, lab = loss_inputs(512, 768, (True, True, True), seed=3, labels="multi", dtype=torch.bfloat16)
    ref_in = [f.float() for f in feats]
    ref_loss, ref_grads = loss_oracle.closed_form_grads(ref_in, lab)
    monkeypatch.setenv("BSC_BWD_MEM_GB", "0.0001")  # 107 KB < 3 pairs x 512 x 512 x 2 B
    loss, grads = _run(ref_in, lab, "tc16")
    assert _lib.last_bwd_mode() == 0
    _compare(loss, grads, ref_loss, ref_grads, TC16_TOL)
    monkeypatch.delenv("BSC_BWD_MEM_GB")
    _run(ref_in, lab, "tc16")
    assert _lib.last_bwd_mode() == 4


@pytest.mark.parametrize("n,d,present,labels,shared,world,normalised", [
    (512, 768, (True, True, True), "multi", 0.5, 2, True),
    (1024, 768, (True, True, True), "arange", 0.5, 4, True),
    (768, 512, (True, False, True), "multi", 0.5, 2, True),      # 384 rows per rank: three row blocks, d = 2 x 256
    (2048, 768, (True, True, True), "multi", 0.5, 8, True),
    (1024, 768, (True, True, True), "multi", 0.5, 2, False),     # row norms 0.5 .. 2 x sqrt(d): 1/|row| in the weights
    (1200, 768, (True, True, True), "multi", 0.5, 4, True),      # 300 rows per rank: row blocks straddle two owners
    (1040, 384, (True, True, True), "multi", 0.5, 8, False),     # 130 rows per rank, d = one 384-wide item
])
@pytest.mark.parametrize("mode", ["scatter", "blocks", "recompute"])
def test_tc16_ranks_replayed_on_one_gpu(n, d, present, labels, shared, world, normalised, mode):
    """The kernels the ranks of a multi-GPU job run, replayed on ONE GPU rank by rank and phase by phase (tests/stripe.py):
    forward statistics of every row stripe, summed as the all-reduce would; then the backward --
      scatter    weight block per UNORDERED pair on the local rows, dA locally, the partial dB of ALL rows stored into the
                 owners' exchange buffers (here W buffers on the one device), fixed-order sum in the finish phase;
      blocks     weight block per ORDERED pair, no exchange;     recompute   the fused sweep, no exchange.
    Loss and gradients of the whole batch against the oracle (loss_func.py:122-162: every rank evaluates the global loss,
    the gradient reaches the local rows; with the scatter strategy through the backward of the autograd-aware
    all_gather, loss_func.py:81-82, whose sum over ranks is exactly this reduce-scatter)."""
    from bsc._lib import BWD_NAMES
    from bsc.loss import _default_ops
    from tests.stripe import replay

    if mode == "recompute" and d not in (256, 512, 768):
        pytest.skip("the recomputing sweep covers d in {256, 512, 768}")
    feats, lab = loss_inputs(n, d, present, seed=n + world, labels=labels, shared=shared, dtype=torch.bfloat16,
                             normalised=normalised)
    ref_in = [None if f is None else f.float() for f in feats]
    ref_loss, ref_grads = loss_oracle.closed_form_grads(ref_in, lab)
    ops = _default_ops()
    dev_feats = [None if f is None else f.cuda() for f in ref_in]
    losses, grads, resolved = replay(ops, world, dev_feats, lab.cuda(), BWD_NAMES[mode])
    torch.cuda.synchronize()
    assert resolved == BWD_CODE[mode]
    assert all(x == losses[0] for x in losses)  # every rank reports the loss of the global batch, same bits
    _compare(losses[0], [None if g is None else g.cpu() for g in grads], ref_loss, ref_grads, TC16_TOL)


def test_strided_rows_and_mixed_dtypes():
    """Inputs as encoders hand them over: a CLS slice hidden[:, 0] (row stride T*D) and a column slice [:, :d].  The
    forward reads them in place; the backward must neither read nor write outside the rows (the gradients come back
    dense).  Mixed dtypes are refused up front."""
    import bsc

    n, d, t = 192, 256, 3
    feats, lab = loss_inputs(n, d, seed=21, labels="multi", dtype=torch.bfloat16)
    ref_in = [f.float() for f in feats]
    ref_loss, ref_grads = loss_oracle.closed_form_grads(ref_in, lab)
    hidden = torch.full((n, t, d), 7.0, device="cuda")
    hidden[:, 0] = ref_in[0].cuda()
    wide = torch.full((n, 2 * d), -3.0, device="cuda")
    wide[:, :d] = ref_in[1].cuda()
    leaves = [hidden.clone().requires_grad_(True), wide.clone().requires_grad_(True),
              ref_in[2].cuda().requires_grad_(True)]
    views = [leaves[0][:, 0], leaves[1][:, :d], leaves[2]]
    assert not views[0].is_contiguous() and not views[1].is_contiguous()
    for precision, tol in (("tc16", TC16_TOL), ("fp32", TC16_TOL)):
        for l in leaves:
            l.grad = None
        loss = bsc.ContrastiveLoss(None, 1 / 0.07, precision=precision)(views[0], views[1], views[2], lab.cuda())
        loss.backward()
        torch.cuda.synchronize()
        got = [leaves[0].grad[:, 0].cpu(), leaves[1].grad[:, :d].cpu(), leaves[2].grad.cpu()]
        _compare(loss.detach().cpu(), got, ref_loss, ref_grads, tol)
        assert leaves[0].grad[:, 1:].abs().max().item() == 0.0 and leaves[1].grad[:, d:].abs().max().item() == 0.0
    with pytest.raises(ValueError, match="one dtype"):
        bsc.ContrastiveLoss(None, 1 / 0.07)(views[0], views[1].bfloat16(), None, lab.cuda())


def test_tc16_unnormalised_rows_of_mixed_scale():
    """Rows whose norms span 0.05 .. 40: the inverse norms live in the fp32 epilogue and in the weight tile."""
    feats, lab = loss_inputs(640, 768, seed=77, labels="multi", normalised=False, dtype=torch.bfloat16)
    g = torch.Generator().manual_seed(5)
    scl = torch.exp(torch.empty(640, 1).uniform_(-3.0, 3.7, generator=g))
    feats = [(f.float() * scl).bfloat16() for f in feats]
    ref_in = [f.float() for f in feats]
    ref_loss, ref_grads = loss_oracle.closed_form_grads(ref_in, lab)
    loss, grads = _run(ref_in, lab, "tc16")
    assert abs(float(loss) - float(ref_loss)) <= TC16_TOL * abs(float(ref_loss))
    for g_, gr in zip(grads, ref_grads):
        # per-row bound: a row's gradient scales with 1/||x_i||, so compare row by row
        err = (g_.double() - gr.double()).abs().amax(dim=1)
        assert (err <= TC16_TOL * gr.abs().amax(dim=1) + 1e-12).all(), (err / gr.abs().amax(dim=1)).max().item()


def test_tc16_config2_4096_three_modalities():
    """BASELINE config 2: image-DNA-text, global batch 4096, bf16 inputs."""
    feats, lab = loss_inputs(4096, 768, seed=7, labels="arange", dtype=torch.bfloat16)
    ref_in = [f.float() for f in feats]
    torch.set_num_threads(os.cpu_count() or 8)
    ref_loss, ref_grads = loss_oracle.contrastive_loss_and_grads(ref_in, lab)
    loss, grads = _run(ref_in, lab, "tc16")
    _compare(loss, grads, ref_loss, ref_grads, TC16_TOL)
    loss_b, grads_b = _run(feats, lab, "tc16")   # bf16 leaves: one extra rounding of the result
    assert abs(float(loss_b) - float(ref_loss)) <= TC16_TOL * abs(float(ref_loss)) + LOSS_FLOOR
    for g, gr in zip(grads_b, ref_grads):
        assert ((g.double() - gr.double()).abs() <= TC16_TOL * gr.abs().max() + _bf16_round_bound(gr)).all()


def _sampled_check(feats, lab, run, rows, tol):
    """Loss and a strided sample of gradient rows against the block-wise fp64 closed form evaluated on the device."""
    ref_in = [None if f is None else f.float() for f in feats]
    ref_loss, ref_rows = loss_oracle.blockwise_loss_and_grad_rows(ref_in, lab, sample_rows=rows, device="cuda")
    loss, grads = run()
    rel_loss = abs(float(loss) - float(ref_loss)) / abs(float(ref_loss))
    assert rel_loss <= tol, (float(loss), float(ref_loss))
    worst = 0.0
    for g, gr in zip(grads, ref_rows):
        if gr is None:
            continue
        # the bound is relative to the largest gradient entry of the WHOLE batch; the sample's own maximum is a lower
        # bound of it, so the check is the stricter one
        err = (g[rows].double() - gr.double()).abs().max().item() / gr.abs().max().item()
        worst = max(worst, err)
        assert err <= tol, err
    return rel_loss, worst


@pytest.mark.parametrize("mode", ["scatter", "recompute"])
def test_config3_32768_against_blockwise_oracle(mode, monkeypatch):
    """BASELINE config 3 on one GPU (global batch 32768, three modalities, bf16-representable inputs held in float32
    so that the returned gradients are not re-rounded): loss and 1536 gradient rows per modality against the fp64 closed
    form of the reference maths, evaluated block-wise (2048 rows x all columns at a time, plain torch fp64 matmuls on
    the device: the reference module itself needs > 100 GB of N x N tensors at this size)."""
    n = 32768
    feats, lab = loss_inputs(n, 768, seed=3, labels="multi", dtype=torch.bfloat16)
    rows = torch.arange(5, n, 21)[:1536]
    monkeypatch.setenv("BSC_BWD_MODE", mode)
    rel_loss, worst = _sampled_check(feats, lab, lambda: _run([f.float() for f in feats], lab, "tc16"), rows, TC16_TOL)
    print(f"N=32768 {mode}: loss rel err {rel_loss:.2e}, gradient max err / max |g| {worst:.2e}")


def test_config3_32768_eight_ranks_replayed_against_blockwise_oracle():
    """Same batch through the 8-rank code path (4096 rows per rank, scatter backward) replayed on one GPU."""
    from bsc._lib import BWD_SCATTER
    from bsc.loss import _default_ops
    from tests.stripe import replay

    n = 32768
    feats, lab = loss_inputs(n, 768, seed=3, labels="multi", dtype=torch.bfloat16)
    rows = torch.arange(5, n, 21)[:1536]
    dev = [f.float().cuda() for f in feats]
    out = {}

    def run():
        losses, grads, mode = replay(_default_ops(), 8, dev, lab.cuda(), BWD_SCATTER)
        torch.cuda.synchronize()
        assert mode == BWD_SCATTER and all(x == losses[0] for x in losses)
        out["loss"] = losses[0]
        return losses[0], [g.cpu() for g in grads]

    rel_loss, worst = _sampled_check(feats, lab, run, rows, TC16_TOL)
    print(f"N=32768 as 8 ranks: loss rel err {rel_loss:.2e}, gradient max err / max |g| {worst:.2e}")


X3_TOL = 1e-5  # north_star: 1e-5 relative in fp32 -- here on the tensor cores (BSC_MODE_TC16X3, precision="fp32x3")
X3_CASES = [
    # n, d, present, labels, shared, normalised
    (512, 768, (True, True, True), "multi", 0.5, True),
    (1000, 768, (True, True, True), "arange", 0.5, True),     # ragged N, peaked softmax (loss ~ 1e-2)
    (384, 256, (True, False, True), "multi", 0.0, True),      # flat softmax, two modalities
    (300, 1024, (True, True, True), "multi", 0.5, False),     # un-normalised rows (norms 0.5 .. 2 sqrt(d)), d = 4 x 256
    (4096, 768, (True, True, True), "multi", 0.5, True),      # BASELINE config 2 size, full float32 inputs
    (4096, 768, (True, True, True), "arange", 0.5, True),     # ... with the reference's default labels
]


def _x3_errors(loss, grads, ref_loss, ref_grads):
    rel_loss = abs(float(loss) - float(ref_loss)) / abs(float(ref_loss))
    rel_grad = max((g.double() - r.double()).abs().max().item() / r.abs().max().item()
                   for g, r in zip(grads, ref_grads) if r is not None)
    return rel_loss, rel_grad


@pytest.mark.parametrize("n,d,present,labels,shared,normalised", X3_CASES)
@pytest.mark.parametrize("mode", ["scatter", "blocks"])
def test_fp32x3_tensor_core_mode_matches_fp64_closed_form(n, d, present, labels, shared, normalised, mode, monkeypatch):
    """Float32 inputs (NOT 16-bit representable) through the split-operand tensor-core mode against the fp64 closed form
    of the reference maths.  With labels that give a loss of order 1 the bound is the plain 1e-5 on the loss and on
    max |gradient error| / max |gradient|; with the reference's default labels (arange: a peaked softmax and a loss of
    order 1e-2, where the reference's own float32 arithmetic is itself several 1e-5 away from the closed form) the
    float32 noise of the reference is allowed on top, as in test_fp32_mode_matches_oracle.  The measured figures are
    printed either way (pytest -s)."""
    from bsc import _lib

    feats, lab = loss_inputs(n, d, present, seed=n + 3, labels=labels, shared=shared, normalised=normalised)
    torch.set_num_threads(os.cpu_count() or 8)
    ref_loss, ref_grads = loss_oracle.closed_form_grads(feats, lab)
    nl, ng = _fp32_noise(feats, lab, ref_loss, ref_grads)
    monkeypatch.setenv("BSC_BWD_MODE", mode)
    loss, grads = _run(feats, lab, "fp32x3")
    assert _lib.last_bwd_mode() == BWD_CODE[mode]
    rel_loss, rel_grad = _x3_errors(loss, grads, ref_loss, ref_grads)
    ref_gmax = max(r.abs().max().item() for r in ref_grads if r is not None)
    print(f"fp32x3 N={n} d={d} {labels} {mode}: loss {float(ref_loss):.5f} rel err {rel_loss:.2e}, gradient max err / "
          f"max|g| {rel_grad:.2e}  (float32 reference port vs the same closed form: {nl / abs(float(ref_loss)):.2e} / "
          f"{max(ng) / ref_gmax:.2e})")
    if labels == "arange":
        _compare(loss, grads, ref_loss, ref_grads, X3_TOL, nl, ng)
    else:
        assert rel_loss <= X3_TOL + LOSS_FLOOR / abs(float(ref_loss)), rel_loss
        assert rel_grad <= X3_TOL, rel_grad


def test_fp32x3_config3_32768_against_blockwise_oracle():
    """The headline batch (N = 32768, d = 768, full float32 inputs) in the fp32x3 mode: loss and 1024 gradient rows per
    modality against the block-wise fp64 closed form.  At this size the K-major gradient GEMMs are cut into 32 chains per
    row block -- 25344 work items, more than one launch's schedule holds: the gradient GEMMs run as several launches."""
    n = 32768
    feats, lab = loss_inputs(n, 768, seed=11, labels="multi")
    rows = torch.arange(3, n, 31)[:1024]
    rel_loss, worst = _sampled_check(feats, lab, lambda: _run(feats, lab, "fp32x3"), rows, 2e-5)
    print(f"fp32x3 N=32768: loss rel err {rel_loss:.2e}, gradient max err / max|g| {worst:.2e}")
    assert rel_loss <= X3_TOL, rel_loss


@pytest.mark.parametrize("world,mode,n", [(2, "scatter", 1024), (4, "scatter", 1024), (4, "blocks", 1024),
                                          (2, "scatter", 4096)])
def test_fp32x3_ranks_replayed_on_one_gpu(world, mode, n):
    """The W-rank job of the fp32x3 mode (rows gathered with plain collectives, scatter backward over the exchange
    buffers) replayed on one GPU: loss bits equal on every rank, loss and gradients of the whole batch at 1e-5.
    n = 4096 on 2 ranks: 2048 local rows = two accumulation chains per scattered GEMM, i.e. two exchange slots per
    source rank."""
    from bsc._lib import BWD_NAMES, MODE_TC16X3
    from bsc.loss import _default_ops
    from tests.stripe import replay

    d = 768
    feats, lab = loss_inputs(n, d, seed=n + world, labels="multi", normalised=False)
    ref_loss, ref_grads = loss_oracle.closed_form_grads(feats, lab)
    losses, grads, resolved = replay(_default_ops(), world, [f.cuda() for f in feats], lab.cuda(), BWD_NAMES[mode],
                                     mode=MODE_TC16X3)
    torch.cuda.synchronize()
    assert resolved == BWD_CODE[mode] and all(x == losses[0] for x in losses)
    rel_loss, rel_grad = _x3_errors(losses[0], [g.cpu() for g in grads], ref_loss, ref_grads)
    print(f"fp32x3 as {world} ranks ({mode}): loss rel err {rel_loss:.2e}, gradient max err / max|g| {rel_grad:.2e}")
    assert rel_loss <= X3_TOL and rel_grad <= X3_TOL


def test_fp32x3_is_what_auto_picks_for_large_float32_batches_and_refuses_the_recomputing_sweep(monkeypatch):
    import bsc
    from bsc._lib import MODE_F32, MODE_TC16X3
    from bsc.loss import resolve_mode

    small, big = [torch.zeros(256, 768)] * 2, [torch.zeros(2048, 768)] * 2
    assert resolve_mode("auto", small) == MODE_F32 and resolve_mode("auto", big) == MODE_TC16X3
    assert resolve_mode("auto", small, n_global=8 * 256) == MODE_TC16X3  # the GLOBAL batch decides
    feats, lab = loss_inputs(256, 768, seed=5, labels="multi")
    monkeypatch.setenv("BSC_BWD_MODE", "recompute")
    with pytest.raises(bsc.BscError, match="stored-weight"):
        _run(feats, lab, "fp32x3")


def test_properties_at_full_size():
    """Size-independent checks at N = 32768: permutation invariance of the batch and run-to-run determinism."""
    n = 32768
    feats, lab = loss_inputs(n, 768, seed=3, labels="multi", dtype=torch.bfloat16)
    loss, grads = _run(feats, lab, "tc16")
    perm = torch.randperm(n, generator=torch.Generator().manual_seed(5))
    loss_p, grads_p = _run([f[perm] for f in feats], lab[perm], "tc16")
    assert abs(float(loss) - float(loss_p)) <= 1e-4 * abs(float(loss))
    for g, gp in zip(grads, grads_p):
        # both sides are rounded to bf16 once: allow two roundings on top of the 2e-3 arithmetic bound
        assert ((g[perm] - gp).abs() <= 2e-3 * g.abs().max() + 2 * _bf16_round_bound(gp)).all()
    assert all(torch.isfinite(g).all() for g in grads)
    # run-to-run determinism: fixed summation orders everywhere, so repeated evaluations give the same bits
    loss2, grads2 = _run(feats, lab, "tc16")
    assert float(loss) == float(loss2)
    assert all(torch.equal(g, g2) for g, g2 in zip(grads, grads2))


def test_error_behaviour():
    import bsc

    x = torch.randn(8, 64, device="cuda")
    with pytest.raises(ValueError, match="Too less element"):
        bsc.ContrastiveLoss(torch.nn.CrossEntropyLoss(), 1 / 0.07)(x, None, None, torch.arange(8).cuda())
    with pytest.raises(NotImplementedError):
        bsc.ContrastiveLoss(torch.nn.CrossEntropyLoss(label_smoothing=0.1), 1 / 0.07)
    with pytest.raises(bsc.BscError):
        bsc.ContrastiveLoss(None, 1 / 0.07)(x.cpu(), x.cpu(), None, torch.arange(8))
    # d not a multiple of 128 is a loud error in the tensor-core mode, not a silent detour
    y = torch.randn(8, 72, device="cuda", dtype=torch.bfloat16)
    with pytest.raises(bsc.BscError, match="BSC_MODE_F32"):
        bsc.ContrastiveLoss(None, 1 / 0.07, precision="tc16")(y, y, None, torch.arange(8).cuda())
    # ... while 'auto' takes the fp32 path for widths the tensor-core kernels do not tile
    la = bsc.ContrastiveLoss(None, 1 / 0.07, precision="auto")(y.requires_grad_(True), y.detach(), None,
                                                               torch.arange(8).cuda())
    la.backward()
    assert torch.isfinite(la) and y.grad is not None
    # a temperature that wants a gradient is refused (the reference's is a constant, train_cl.py:190)
    with pytest.raises(NotImplementedError, match="constant logit_scale"):
        bsc.ClipLoss()(x, x, None, torch.arange(8).cuda(), torch.tensor(14.0, device="cuda", requires_grad=True))


def test_cliploss_single_rank_output_dict():
    import bsc

    feats, lab = loss_inputs(96, 64, seed=9, labels="multi")
    ref_loss, _ = loss_oracle.closed_form_grads(feats, lab)
    crit = bsc.ClipLoss(rank=0, world_size=1, criterion=torch.nn.CrossEntropyLoss())
    out = crit(feats[0].cuda(), feats[1].cuda(), feats[2].cuda(), lab.cuda(), 1 / 0.07, output_dict=True)
    assert abs(float(out["contrastive_loss"]) - float(ref_loss)) <= FP32_TOL * abs(float(ref_loss))


@pytest.mark.parametrize("n,classes", [(1, 1), (300, 50), (5000, 5000), (4097, 3), (20000, 700)])
def test_label_stats_match_label_matrix(n, classes):
    """r / rep / next / rmax of bsc_label_stats vs construct_label_metrix (loss_func.py:18-21), bit-exact."""
    from bsc.ops import CudaOps

    g = torch.Generator().manual_seed(n)
    lab = torch.randint(-(2 ** 40), 2 ** 40, (classes,), generator=g)[torch.randint(0, classes, (n,), generator=g)]
    if classes == n:
        lab = torch.arange(n) * 7919 - 3
    r, rep, nxt, rmax = CudaOps().label_stats(lab.cuda())
    ln = lab.numpy()
    order = np.argsort(ln, kind="stable")
    want_r = np.empty(n, np.float32)
    want_rep = np.empty(n, np.int32)
    want_next = np.full(n, -1, np.int32)
    start = 0
    while start < n:
        end = start
        while end < n and ln[order[end]] == ln[order[start]]:
            end += 1
        members = order[start:end]
        want_r[members] = end - start
        want_rep[members] = members[0]
        want_next[members[:-1]] = members[1:]
        start = end
    assert np.array_equal(r.cpu().numpy(), want_r)
    assert np.array_equal(rep.cpu().numpy(), want_rep)
    assert np.array_equal(nxt.cpu().numpy(), want_next)
    assert float(rmax) == want_r.max()


def test_synthetic_train_step_with_fused_loss():
    """Caller side (train_epoch.py:29-40 + simple_clip.py:27-50) with small random-init encoders: the loss drives LoRA
    and head gradients only, base weights stay frozen, and the loss value equals the oracle's on the same embeddings."""
    import bsc
    from bsc import harness

    torch.manual_seed(1)
    model = harness.build_synthetic_model(vit_layers=1, dna_layers=1, text_layers=1).cuda()
    params = [p for p in model.parameters() if p.requires_grad]
    before = [p.detach().clone() for p in params]
    frozen = model.image_encoder.vit.encoder.layer[0].attention.attention.query.base.weight
    frozen_before = frozen.detach().clone()
    opt = torch.optim.SGD(params, lr=0.1)
    batch = harness.synthetic_batch(48, "cuda", seed=2)
    crit = bsc.ContrastiveLoss(torch.nn.CrossEntropyLoss(), 1 / 0.07, precision="fp32")
    loss = harness.train_step(model, crit, opt, batch, autocast_dtype=None)
    assert torch.isfinite(loss)
    assert torch.equal(frozen, frozen_before) and frozen.grad is None
    assert any(not torch.equal(p, b) for p, b in zip(params, before))
    with torch.no_grad():
        outs = model(*batch[:3])
    ref, _ = loss_oracle.closed_form_grads([o.float().cpu() for o in outs], batch[3].cpu())
    got = crit(*[o.detach().requires_grad_(True) for o in outs], batch[3])
    assert abs(float(got) - float(ref)) <= FP32_TOL * abs(float(ref)) + LOSS_FLOOR


def test_backward_on_the_autograd_worker_thread_with_a_warm_allocator():
    """torch runs the backward on its own worker thread.  When every buffer comes out of the caching allocator, the first
    CUDA call of that thread is this library's tensor-map encoding (a DRIVER call): the context must be bound by the
    library itself (CUDA_ERROR_INVALID_CONTEXT in the 8-GPU train step otherwise).  Fresh process, warm allocator."""
    import subprocess
    import sys

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    code = (
        "import sys, torch\n"
        f"sys.path.insert(0, {root!r}); sys.path.insert(0, {os.path.join(root, 'bioscan-clip_b200')!r})\n"
        "import bsc\n"
        "warm = torch.empty(1 << 30, dtype=torch.uint8, device='cuda'); del warm\n"
        "g = torch.Generator(device='cuda').manual_seed(0)\n"
        "x = [torch.nn.functional.normalize(torch.randn(512, 768, device='cuda', generator=g), dim=1).bfloat16()"
        ".requires_grad_(True) for _ in range(3)]\n"
        "loss = bsc.ContrastiveLoss(None, 1 / 0.07, precision='tc16')(x[0], x[1], x[2], torch.arange(512, device='cuda'))\n"
        "loss.backward(); torch.cuda.synchronize()\n"
        "assert all(torch.isfinite(t.grad).all() for t in x)\n"
        "print('ok', float(loss))\n")
    r = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0 and "ok" in r.stdout, r.stderr[-1500:]
