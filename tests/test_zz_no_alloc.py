"""bsc_label_stats_static (the entry point that allocates nothing) against bsc_label_stats, through the C ABI.

Collected LAST on purpose (file name): a failure here must not hide other tests.
The results are integers and integer-valued floats produced by the same kernels, so the comparison is bit-exact, as in
the other label-statistics tests.  The static scratch array is shared by every call and keeps the previous call's
contents: the test calls twice in a row, with DIFFERENT labels of different sizes, so that stale table slots / counts
of the first call would show up in the second.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_N = 65536  # BSC_LABEL_STATS_STATIC_MAX_N in include/bsc_hot.h


def _outputs(n, dev):
    return (torch.full((max(n, 1),), -7.0, dtype=torch.float32, device=dev),
            torch.full((max(n, 1),), -7, dtype=torch.int32, device=dev),
            torch.full((max(n, 1),), -7, dtype=torch.int32, device=dev),
            torch.full((1,), -7.0, dtype=torch.float32, device=dev))


def _existing(L, labels):
    n, dev = labels.numel(), labels.device
    r, rep, nxt, rmax = _outputs(n, dev)
    nbytes = ctypes.c_size_t(0)
    assert L.bsc_label_stats_workspace_bytes(n, ctypes.byref(nbytes)) == 0
    ws = torch.empty(int(nbytes.value), dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    rc = L.bsc_label_stats(labels.data_ptr(), n, r.data_ptr(), rep.data_ptr(), nxt.data_ptr(), rmax.data_ptr(),
                           ws.data_ptr(), ws.numel(), stream)
    assert rc == 0, L.bsc_last_error().decode()
    torch.cuda.synchronize(dev)
    return r[:n], rep[:n], nxt[:n], rmax


def _static(L, labels, outs):
    """The call under test: inputs and outputs only; the output tensors are allocated by the caller BEFORE the call."""
    n, dev = labels.numel(), labels.device
    r, rep, nxt, rmax = outs
    stream = torch.cuda.current_stream(dev).cuda_stream
    before = torch.cuda.memory_stats(dev)["allocation.all.allocated"]
    rc = L.bsc_label_stats_static(labels.data_ptr(), n, r.data_ptr(), rep.data_ptr(), nxt.data_ptr(), rmax.data_ptr(),
                                  stream)
    assert rc == 0, L.bsc_last_error().decode()
    assert torch.cuda.memory_stats(dev)["allocation.all.allocated"] == before  # nothing went through torch's allocator
    torch.cuda.synchronize(dev)
    return r[:n], rep[:n], nxt[:n], rmax


def _labels(kind, n, dev, seed):
    g = torch.Generator(device="cpu").manual_seed(seed)
    if kind == "arange":  # the reference's default labels (util/dataset.py:147): every row alone in its class
        return torch.arange(n, dtype=torch.int64).to(dev)
    if kind == "classes":  # many rows per class, 64-bit label values
        return (torch.randint(0, max(n // 7, 1), (n,), generator=g, dtype=torch.int64) * 2654435761 - (1 << 40)).to(dev)
    return torch.full((n,), 42, dtype=torch.int64).to(dev)  # one class


def test_static_label_stats_matches_bsc_label_stats_two_calls_in_a_row():
    from bsc import _lib

    dev = torch.device("cuda", 0)
    _lib.require_device(dev)
    L = _lib.lib()
    with torch.cuda.device(dev):
        # two calls in a row on the shared scratch: a large dense-class problem first, then a smaller one whose table
        # is a prefix of the first call's, then the maximum size, then the smallest ones
        for step, (kind, n) in enumerate([("classes", 32768), ("arange", 4099), ("classes", MAX_N), ("one", 257),
                                          ("classes", 1)]):
            labels = _labels(kind, n, dev, seed=100 + step)
            want = _existing(L, labels)
            outs = _outputs(n, dev)
            got = _static(L, labels, outs)
            for name, a, b in zip(("r", "rep", "next", "rmax"), got, want):
                assert torch.equal(a, b), f"call {step} ({kind}, n={n}): {name} differs from bsc_label_stats"
            assert float(got[3]) == float(got[0].max())


def test_static_label_stats_rejects_sizes_beyond_the_static_scratch():
    from bsc import _lib

    dev = torch.device("cuda", 0)
    _lib.require_device(dev)
    L = _lib.lib()
    n = MAX_N + 1
    with torch.cuda.device(dev):
        labels = torch.arange(n, dtype=torch.int64, device=dev)
        r, rep, nxt, rmax = _outputs(n, dev)
        rc = L.bsc_label_stats_static(labels.data_ptr(), n, r.data_ptr(), rep.data_ptr(), nxt.data_ptr(),
                                      rmax.data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
        torch.cuda.synchronize(dev)
    assert rc == -2  # BSC_EUNSUPPORTED
    msg = L.bsc_last_error().decode()
    assert "65536" in msg and "bsc_label_stats" in msg
    assert float(r[0]) == -7.0 and float(rmax[0]) == -7.0  # a rejected call touches nothing
