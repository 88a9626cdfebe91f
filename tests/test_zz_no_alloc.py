"""GPU: dq4ml_dq_filter_noalloc, the DQ filter for callers that are not allowed to allocate GPU memory.

Some embedders hand the whole GPU to their own memory pool before loading this library, or call the filter from a
path where cudaMalloc's device-wide stall is unacceptable; for them the library offers an entry point that keeps its
scratch in a fixed array inside the library image and never calls an allocator.  This file checks, through the C ABI,
that it returns exactly what the regular dq4ml_dq_filter returns -- bit for bit, the tolerance every filter test here
uses -- on two calls in a row (the second one smaller and with different data, so words the first call left in the
shared array would show up), and that it refuses inputs beyond its stated row limit.

The file name sorts last on purpose: a run that stops at the first failure has already been through everything else.
It needs nothing from the other tests and passes on its own in a fresh process too.
"""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

MAX_ROWS = 1 << 26  # DQ4ML_NOALLOC_MAX_ROWS in include/dq4ml.h


def _call_noalloc(L, guest, price, stream=None):
    n = guest.numel()
    go, po = torch.empty_like(guest), torch.empty_like(price)
    n_out, n1 = C.c_int64(-1), C.c_int64(-1)
    rc = L.dq4ml_dq_filter_noalloc(stream, guest.data_ptr(), price.data_ptr(), n, None, go.data_ptr(), po.data_ptr(),
                                   C.byref(n_out), C.byref(n1))
    return rc, go, po, int(n_out.value), int(n1.value)


def _call_regular(pkg, ctx, guest, price):
    go, po, n1 = pkg.dq_filter(ctx, guest, price)
    return go, po, n1


def test_noalloc_filter_matches_regular_filter_on_two_calls_in_a_row(pkg, ctx):
    L = pkg.capi.lib()
    torch.cuda.set_device(0)
    # two different inputs: a large one with the default anomaly rates, then a smaller one where half the rows go
    first = ctx.synth_generate(5_000_003, d=1, row0=0, p_low=0.01, p_high=0.01)[:2]
    second = ctx.synth_generate(1_234_567, d=1, row0=77_000_000, p_low=0.25, p_high=0.25)[:2]
    for guest, price in (first, second):
        ref_g, ref_p, ref_n1 = _call_regular(pkg, ctx, guest, price)
        rc, go, po, m, n1 = _call_noalloc(L, guest, price)
        assert rc == 0, L.dq4ml_last_error()
        assert (m, n1) == (ref_g.shape[0], ref_n1)
        assert torch.equal(go[:m], ref_g) and torch.equal(po[:m].view(torch.int64), ref_p.view(torch.int64))
    # same call again on a caller-owned stream: still identical
    s = torch.cuda.Stream()
    guest, price = first
    ref_g, ref_p, ref_n1 = _call_regular(pkg, ctx, guest, price)
    torch.cuda.synchronize()
    rc, go, po, m, n1 = _call_noalloc(L, guest, price, stream=s.cuda_stream)
    assert rc == 0 and (m, n1) == (ref_g.shape[0], ref_n1)
    assert torch.equal(go[:m], ref_g) and torch.equal(po[:m].view(torch.int64), ref_p.view(torch.int64))


def test_noalloc_filter_edge_sizes_and_limit(pkg, ctx):
    L = pkg.capi.lib()
    torch.cuda.set_device(0)
    # around the tile boundaries of both kernels (2048 and 4096 rows), and inputs of several hundred tiles
    for n in (0, 1, 2047, 2048, 2049, 4095, 4096, 4097, 4096 * 257 + 1, 2048 * 1031 + 5):
        guest = torch.arange(n, dtype=torch.int32, device="cuda") % 36
        price = (torch.arange(n, dtype=torch.float64, device="cuda") * 7.5) % 160.0
        torch.cuda.synchronize()  # the ctx stream is non-blocking: it does not wait for torch's stream to fill the inputs
        ref_g, ref_p, ref_n1 = _call_regular(pkg, ctx, guest, price)
        rc, go, po, m, n1 = _call_noalloc(L, guest, price)
        assert rc == 0 and (m, n1) == (ref_g.shape[0], ref_n1)
        assert torch.equal(go[:m], ref_g) and torch.equal(po[:m], ref_p)
    # one row past the limit is refused before anything is launched; the buffers are never touched
    g = torch.zeros(16, dtype=torch.int32, device="cuda")
    p = torch.zeros(16, dtype=torch.float64, device="cuda")
    n_out = C.c_int64(-1)
    rc = L.dq4ml_dq_filter_noalloc(None, g.data_ptr(), p.data_ptr(), MAX_ROWS + 1, None, g.data_ptr() + 0, p.data_ptr(),
                                   C.byref(n_out), None)
    assert rc == pkg.capi.ERR_UNSUPPORTED
    assert b"DQ4ML_NOALLOC_MAX_ROWS" in L.dq4ml_last_error()


@pytest.mark.parametrize("n", [MAX_ROWS, MAX_ROWS - 4095])
def test_noalloc_filter_at_the_row_limit(pkg, ctx, n):
    """The largest accepted input uses the last look-back word of the built-in array (n = 2^26 is exactly 32768 tiles of 2048 rows);
    2^26 - 4095 ends on a tile holding a single row."""
    L = pkg.capi.lib()
    torch.cuda.set_device(0)
    guest, price = ctx.synth_generate(n, d=1, row0=3, p_low=0.05, p_high=0.05)[:2]
    ref_g, ref_p, ref_n1 = _call_regular(pkg, ctx, guest, price)
    rc, go, po, m, n1 = _call_noalloc(L, guest, price)
    assert rc == 0, L.dq4ml_last_error()
    assert (m, n1) == (ref_g.shape[0], ref_n1)
    assert torch.equal(go[:m], ref_g) and torch.equal(po[:m].view(torch.int64), ref_p.view(torch.int64))
    # and once more: the words the first call left behind must not leak into the second
    rc, go2, po2, m2, n12 = _call_noalloc(L, guest, price)
    assert rc == 0 and (m2, n12) == (m, n1)
    assert torch.equal(go2[:m], ref_g) and torch.equal(po2[:m].view(torch.int64), ref_p.view(torch.int64))


def test_noalloc_filter_unaligned_slices(pkg, ctx):
    """The kernel uses scalar loads: a column that starts in the middle of a 16-byte line is as good as any other."""
    L = pkg.capi.lib()
    torch.cuda.set_device(0)
    guest, price = ctx.synth_generate(100_003, d=1, row0=11, p_low=0.2, p_high=0.2)[:2]
    g1, p1 = guest[1:], price[1:]  # 4- and 8-byte aligned only
    ref_g, ref_p, ref_n1 = _call_regular(pkg, ctx, g1.clone(), p1.clone())
    rc, go, po, m, n1 = _call_noalloc(L, g1, p1)
    assert rc == 0, L.dq4ml_last_error()
    assert (m, n1) == (ref_g.shape[0], ref_n1)
    assert torch.equal(go[:m], ref_g) and torch.equal(po[:m].view(torch.int64), ref_p.view(torch.int64))
