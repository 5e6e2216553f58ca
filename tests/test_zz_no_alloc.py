"""sgnn_layer_norm_bwd_static (include/sgnn_b200.h): the LayerNorm backward whose scratch is a static
__device__ array inside the library, for callers that cannot allocate.  Called through the C ABI and compared
with sgnn_layer_norm_bwd (caller workspace) on two calls in a row; the two entries launch the same kernels on
the same data, so the bar is bit equality (the LayerNorm tests' 2e-5 tolerance is implied).
Collected last on purpose."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _inputs(rows, width, seed):
    g = torch.Generator().manual_seed(seed)
    x = (3.0 * torch.randn(rows, width, generator=g) + 1.5).cuda()
    dy = torch.randn(rows, width, generator=g).cuda()
    gamma = torch.randn(width, generator=g).cuda()
    mean = x.mean(1)
    rstd = 1.0 / torch.sqrt(x.var(1, unbiased=False) + 1e-5)
    return dy, x, mean.contiguous(), rstd.contiguous(), gamma


def _workspace_entry(L, dy, x, mean, rstd, gamma):
    n, w = x.shape
    dx = torch.empty_like(x)
    dgamma, dbeta = torch.empty_like(gamma), torch.empty_like(gamma)
    nbytes = L.lib().sgnn_layer_norm_bwd_workspace_bytes(w)
    ws = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    L.check(L.lib().sgnn_layer_norm_bwd(L.ptr(dy), L.ptr(x), L.ptr(mean), L.ptr(rstd), L.ptr(gamma), C.c_long(n), w,
                                        L.ptr(dx), L.ptr(dgamma), L.ptr(dbeta), L.ptr(ws), C.c_size_t(nbytes),
                                        L.stream_ptr()), launches=2)
    return dx, dgamma, dbeta


def _static_entry(L, dy, x, mean, rstd, gamma, out):
    n, w = x.shape
    dx, dgamma, dbeta = out
    L.check(L.lib().sgnn_layer_norm_bwd_static(L.ptr(dy), L.ptr(x), L.ptr(mean), L.ptr(rstd), L.ptr(gamma),
                                               C.c_long(n), w, L.ptr(dx), L.ptr(dgamma), L.ptr(dbeta),
                                               L.stream_ptr()), launches=2)


def test_static_scratch_layer_norm_bwd_matches_workspace_entry():
    from sgnn_b200 import _lib as L
    # two calls in a row with different shapes: the second must not see the first call's partials
    cases = [(20011, 128, 1), (777, 64, 2), (5, 32, 3), (300000, 64, 4)]
    wants, outs, args = [], [], []
    for rows, width, seed in cases:
        a = _inputs(rows, width, seed)
        args.append(a)
        wants.append(_workspace_entry(L, *a))
        outs.append((torch.empty_like(a[1]), torch.empty_like(a[4]), torch.empty_like(a[4])))
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    for a, out in zip(args, outs):               # back to back on one stream, outputs preallocated
        _static_entry(L, *a, out)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    for (rows, width, _), want, got in zip(cases, wants, outs):
        for w_, g_, name in zip(want, got, ["dx", "dgamma", "dbeta"]):
            assert torch.isfinite(g_).all(), (rows, width, name)
            assert torch.equal(w_, g_), (rows, width, name)
    # same call twice: identical bits (nothing left over from the previous contents matters)
    again = (torch.empty_like(args[0][1]), torch.empty_like(args[0][4]), torch.empty_like(args[0][4]))
    _static_entry(L, *args[0], again)
    torch.cuda.synchronize()
    assert all(torch.equal(p, q) for p, q in zip(again, outs[0]))


def test_static_scratch_rejects_sizes_beyond_the_maximum():
    from sgnn_b200 import _lib as L
    a = _inputs(8, 256, 9)
    out = (torch.empty_like(a[1]), torch.empty_like(a[4]), torch.empty_like(a[4]))
    with pytest.raises(L.SgnnError, match="exceeds the static scratch"):
        _static_entry(L, *a, out)
    a = _inputs(8, 96, 9)
    out = (torch.empty_like(a[1]), torch.empty_like(a[4]), torch.empty_like(a[4]))
    with pytest.raises(L.SgnnError, match="width must be"):
        _static_entry(L, *a, out)
