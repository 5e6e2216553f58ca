"""The C entries of the hand-written training step, each called through the C ABI on its own and compared with
torch in fp64: sgnn_csr_by_source, sgnn_train_pack + sgnn_train_mlp_fwd / _bwd (hi/lo-split and plain TF32),
sgnn_layer_norm_bwd_rows (gathered rows, device-side row count), sgnn_linear_wgrad_ex (column blocks, scale),
sgnn_linear_wgrad_tc (batched 64 x 64 products on tcgen05, MN-major operands).
Reference semantics: the autograd of build_mlp + LayerNorm (sgnn/single_scale/graph_network.py:21-58,139-147) and
of PyG's propagate gathers (graph_network.py:173)."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
H = 64


def _lib():
    from sgnn_b200 import _lib as L
    lib = L.lib()
    vp, lg, it, fl, sz = C.c_void_p, C.c_long, C.c_int, C.c_float, C.c_size_t
    lib.sgnn_train_pack.argtypes = [vp, it, vp, vp, vp]
    lib.sgnn_train_mlp_fwd.argtypes = [vp, vp]
    lib.sgnn_train_mlp_bwd.argtypes = [vp, vp]
    lib.sgnn_csr_by_source.argtypes = [vp, vp, it, it, vp, vp, vp, sz, vp]
    lib.sgnn_layer_norm_bwd_rows.argtypes = [vp, vp, vp, vp, vp, vp, lg, vp, it, vp, vp, vp, vp, sz, vp]
    lib.sgnn_linear_wgrad_ex.argtypes = [vp, lg, vp, lg, lg, vp, it, it, vp, lg, fl, vp, vp, sz, vp]
    return L, lib


def _images(weights):
    """[(weight [out, in], transposed?)] -> (hi, lo, offsets) through sgnn_train_pack"""
    from sgnn_b200 import train_engine as TE
    L, lib = _lib()
    im = TE._Images()
    offs = [im.transposed(w) if t else im.plain(w) for w, t in weights]
    hi = torch.zeros(im.floats, device="cuda")
    lo = torch.zeros(im.floats, device="cuda")
    arr = (TE.PackDesc * len(im.descs))(*im.descs)
    descs = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8).cuda()
    L.check(lib.sgnn_train_pack(descs.data_ptr(), len(im.descs), hi.data_ptr(), lo.data_ptr(), L.stream_ptr().value))
    return hi, lo, offs


def test_csr_by_source_equals_stable_sort():
    L, lib = _lib()
    g = torch.Generator().manual_seed(0)
    n, cap = 5000, 40000
    for e in (0, 1, 777, 31234):
        src = torch.randint(0, n, (cap,), generator=g, dtype=torch.int32).cuda()
        cnt = torch.tensor([e], dtype=torch.int32, device="cuda")
        rp = torch.empty(n + 1, dtype=torch.int32, device="cuda")
        perm = torch.full((cap,), -1, dtype=torch.int32, device="cuda")
        nb = lib.sgnn_csr_by_source_workspace_bytes(n)
        ws = torch.empty(nb, dtype=torch.uint8, device="cuda")
        L.check(lib.sgnn_csr_by_source(src.data_ptr(), cnt.data_ptr(), cap, n, rp.data_ptr(), perm.data_ptr(),
                                       ws.data_ptr(), nb, L.stream_ptr().value))
        s = src[:e].long()
        want_perm = torch.sort(s, stable=True).indices.int()
        want_rp = torch.zeros(n + 1, dtype=torch.int64, device="cuda")
        want_rp[1:] = torch.cumsum(torch.bincount(s, minlength=n), 0)
        assert torch.equal(rp.long(), want_rp), e
        assert torch.equal(perm[:e], want_perm), e


@pytest.mark.parametrize("split,tol", [(1, 2e-6), (0, 4e-3)])
@pytest.mark.parametrize("rows", [1, 129, 5000])
def test_mlp_forward_and_backward_blocks(split, tol, rows):
    """Linear-ReLU-Linear-LayerNorm with gathered first-layer terms and a residual (the edge / node blocks), then the
    backward chain dpre = (dy W2) * (h1 > 0), dx = dpre W1 with accumulate and residual routing."""
    from sgnn_b200 import train_engine as TE
    L, lib = _lib()
    g = torch.Generator().manual_seed(rows + split)
    cap, n_nodes = rows + 300, 97
    f = lambda *s: torch.randn(*s, generator=g).cuda()      # noqa: E731
    a0, w1, b1, w2, b2, gam, bet = f(cap, H), f(H, H) * 0.2, f(H), f(H, H) * 0.2, f(H), f(H), f(H)
    pq, res = f(n_nodes, 2 * H), f(cap, H)
    tgt = torch.randint(0, n_nodes, (cap,), generator=g, dtype=torch.int32).cuda()
    src = torch.randint(0, n_nodes, (cap,), generator=g, dtype=torch.int32).cuda()
    cnt = torch.tensor([rows], dtype=torch.int32, device="cuda")
    hi, lo, (o1, o2, o2t, o1t) = _images([(w1, False), (w2, False), (w2, True), (w1, True)])
    img = lambda t, o: t.data_ptr() + 4 * o                 # noqa: E731
    h1, y2, y = (torch.full((cap, H), 7.0, device="cuda") for _ in range(3))
    mean, rstd = torch.zeros(cap, device="cuda"), torch.zeros(cap, device="cuda")
    p = TE.TrainFwd()
    p.max_rows, p.n_rows_dev, p.prologue, p.a0, p.ld0, p.k0, p.scale0 = cap, cnt.data_ptr(), 0, a0.data_ptr(), H, H, 2.0
    p.src, p.tgt, p.add_p, p.add_q, p.ld_add = src.data_ptr(), tgt.data_ptr(), pq.data_ptr(), pq.data_ptr() + 4 * H, 2 * H
    p.kpad, p.hidden, p.out, p.n_linear, p.has_ln, p.split = H, H, H, 2, 1, split
    p.w1_hi, p.w1_lo, p.w2_hi, p.w2_lo = img(hi, o1), img(lo, o1), img(hi, o2), img(lo, o2)
    p.b1, p.b1_cols, p.b2, p.gamma, p.beta = b1.data_ptr(), H, b2.data_ptr(), gam.data_ptr(), bet.data_ptr()
    p.res, p.ld_res, p.h1, p.y2, p.mean, p.rstd, p.y, p.ld_y = (res.data_ptr(), H, h1.data_ptr(), y2.data_ptr(),
                                                               mean.data_ptr(), rstd.data_ptr(), y.data_ptr(), H)
    L.check(lib.sgnn_train_mlp_fwd(C.addressof(p), L.stream_ptr().value))
    d = lambda t: t[:rows].double()                         # noqa: E731
    pre = 2.0 * d(a0) @ w1.double().t() + b1.double() + pq.double()[tgt[:rows].long(), :H] + pq.double()[src[:rows].long(), H:]
    wh1 = pre.clamp(min=0)
    wy2 = wh1 @ w2.double().t() + b2.double()
    wy = torch.nn.functional.layer_norm(wy2, (H,), gam.double(), bet.double(), 1e-5) + d(res)
    scale = lambda t: max(1.0, t.abs().max().item())        # noqa: E731
    assert (d(h1) - wh1).abs().max().item() <= tol * scale(wh1)
    assert (d(y2) - wy2).abs().max().item() <= tol * scale(wy2)
    assert (d(y) - wy).abs().max().item() <= 4 * tol * scale(wy)
    assert (d(mean) - wy2.mean(1)).abs().max().item() <= tol * scale(wy2)
    assert (d(rstd) - 1.0 / torch.sqrt(wy2.var(1, unbiased=False) + 1e-5)).abs().max().item() <= 4 * tol
    assert torch.all(h1[rows:] == 7.0) and torch.all(y[rows:] == 7.0)       # rows past the device-side count untouched

    # backward chain from a gradient of the second Linear's output
    dy = f(cap, H)
    dpre = torch.full((cap, H), 7.0, device="cuda")
    dx0, add1 = f(cap, H), f(cap, H)
    dx0_before = dx0.clone()
    q = TE.TrainBwd()
    q.max_rows, q.n_rows_dev, q.dy, q.ld_dy, q.k_dy, q.n_linear, q.hidden = cap, cnt.data_ptr(), dy.data_ptr(), H, H, 2, H
    q.h1, q.dpre, q.w2t_hi, q.w2t_lo = h1.data_ptr(), dpre.data_ptr(), img(hi, o2t), img(lo, o2t)
    q.want_dx, q.kx, q.w1t_hi, q.w1t_lo = 1, H, img(hi, o1t), img(lo, o1t)
    q.dx0, q.ld_dx0, q.k0, q.acc0, q.split = dx0.data_ptr(), H, H, 1, split
    L.check(lib.sgnn_train_mlp_bwd(C.addressof(q), L.stream_ptr().value))
    wdpre = (d(dy) @ w2.double()) * (d(h1) > 0)
    wdx = d(dx0_before) + wdpre @ w1.double()
    assert (d(dpre) - wdpre).abs().max().item() <= tol * scale(wdpre)
    assert (d(dx0) - wdx).abs().max().item() <= 2 * tol * scale(wdx)
    assert torch.all(dpre[rows:] == 7.0) and torch.equal(dx0[rows:], dx0_before[rows:])


def test_layer_norm_bwd_rows_and_wgrad_ex():
    L, lib = _lib()
    g = torch.Generator().manual_seed(3)
    rows, cap, n_nodes = 4321, 6000, 513
    f = lambda *s: torch.randn(*s, generator=g).cuda()      # noqa: E731
    x, dagg, gam = 2.0 * f(cap, H) + 0.5, f(n_nodes, H), f(H)
    idx = torch.randint(0, n_nodes, (cap,), generator=g, dtype=torch.int32).cuda()
    cnt = torch.tensor([rows], dtype=torch.int32, device="cuda")
    mean = x.mean(1).contiguous()
    rstd = (1.0 / torch.sqrt(x.var(1, unbiased=False) + 1e-5)).contiguous()
    dx = torch.zeros(cap, H, device="cuda")
    dgam, dbet = torch.zeros(H, device="cuda"), torch.zeros(H, device="cuda")
    nb = lib.sgnn_layer_norm_bwd_workspace_bytes(H)
    ws = torch.empty(nb, dtype=torch.uint8, device="cuda")
    L.check(lib.sgnn_layer_norm_bwd_rows(dagg.data_ptr(), idx.data_ptr(), x.data_ptr(), mean.data_ptr(), rstd.data_ptr(),
                                         gam.data_ptr(), cap, cnt.data_ptr(), H, dx.data_ptr(), dgam.data_ptr(),
                                         dbet.data_ptr(), ws.data_ptr(), nb, L.stream_ptr().value))
    x64 = x[:rows].double().requires_grad_()
    g64 = gam.double().requires_grad_()
    b64 = torch.zeros(H, dtype=torch.float64, device="cuda", requires_grad=True)
    up = dagg.double()[idx[:rows].long()]
    (torch.nn.functional.layer_norm(x64, (H,), g64, b64, 1e-5) * up).sum().backward()
    assert (dx[:rows].double() - x64.grad).abs().max().item() <= 2e-5 * x64.grad.abs().max().item()
    assert (dgam.double() - g64.grad).abs().max().item() <= 2e-5 * g64.grad.abs().max().item()
    assert (dbet.double() - b64.grad).abs().max().item() <= 2e-5 * b64.grad.abs().max().item()
    assert torch.all(dx[rows:] == 0)

    # weight gradient of a column block: dW[:, 64:128] = 4 * dy[:, :64]^T xx with dy a column block of a wider matrix
    dy_wide, xx = f(cap, 2 * H), f(cap, H)
    dw = torch.full((H, 3 * H), 5.0, device="cuda")
    db = torch.zeros(H, device="cuda")
    nbw = lib.sgnn_linear_wgrad_workspace_bytes(C.c_long(cap), H, H)
    wsw = torch.empty(nbw, dtype=torch.uint8, device="cuda")
    L.check(lib.sgnn_linear_wgrad_ex(dy_wide.data_ptr(), 2 * H, xx.data_ptr(), H, cap, cnt.data_ptr(), H, H,
                                     dw.data_ptr() + 4 * H, 3 * H, 4.0, db.data_ptr(), wsw.data_ptr(), nbw,
                                     L.stream_ptr().value))
    want = 4.0 * dy_wide[:rows, :H].double().t() @ xx[:rows].double()
    assert (dw[:, H:2 * H].double() - want).abs().max().item() <= 1e-5 * want.abs().max().item()
    assert (db.double() - dy_wide[:rows, :H].double().sum(0)).abs().max().item() <= 1e-4
    assert torch.all(dw[:, :H] == 5.0) and torch.all(dw[:, 2 * H:] == 5.0)


@pytest.mark.parametrize("split,tol", [(1, 3e-6), (0, 3e-3)])
@pytest.mark.parametrize("rows,live", [(5000, 4321), (96, 96), (40, 0), (70000, 65537)])
def test_linear_wgrad_tc_batch(split, tol, rows, live):
    """Three 64 x 64 products over one row set in one tcgen05 launch: column-block operands, a scaled result written
    into a column block of a wider gradient, a missing bias gradient, the row count read on the device, rows past
    it (stale, here NaN) never read into the sums; bit-repeatable."""
    from sgnn_b200 import train_engine as TE
    L, lib = _lib()
    torch.manual_seed(rows + split)
    wide = torch.randn(rows, 2 * H, device="cuda")
    xs = [torch.randn(rows, H, device="cuda") for _ in range(2)]
    wide[live:] = float("nan")
    for x in xs:
        x[live:] = float("nan")
    live_dev = torch.tensor([live], dtype=torch.int32, device="cuda")
    ws = torch.empty(lib.sgnn_linear_wgrad_tc_workspace_bytes(3, rows), dtype=torch.uint8, device="cuda")

    def call():
        dw = torch.full((H, 3 * H), 7.0, device="cuda")
        dbs = [torch.full((H,), 7.0, device="cuda") for _ in range(2)]
        b = TE.WgradBatch()
        b.n_items, b.split, b.max_rows, b.n_rows_dev = 3, split, rows, live_dev.data_ptr()
        b.workspace, b.workspace_bytes = ws.data_ptr(), ws.numel()
        spec = [(wide[:, :H], xs[0], 0, 1.0, dbs[0]), (wide[:, H:], xs[0], H, 0.25, None), (wide[:, H:], xs[1], 2 * H, 1.0, dbs[1])]
        for it, (dy, x, col0, scale, db) in zip(b.item, spec):
            it.dy, it.ld_dy, it.x, it.ld_x = dy.data_ptr(), dy.stride(0), x.data_ptr(), x.stride(0)
            it.dw, it.ld_dw, it.scale, it.db = dw.data_ptr() + 4 * col0, dw.stride(0), scale, None if db is None else db.data_ptr()
        L.check(lib.sgnn_linear_wgrad_tc(C.byref(b), L.stream_ptr()))
        torch.cuda.synchronize()
        return dw, dbs, spec

    dw, dbs, spec = call()
    for dy, x, col0, scale, db in spec:
        want = scale * (dy[:live].double().T @ x[:live].double())
        mag = float((dy[:live].double().abs().T @ x[:live].double().abs()).max()) if live else 0.0
        assert float((dw[:, col0:col0 + H].double() - want).abs().max()) <= tol * scale * mag + 1e-30
        if db is not None:
            wb = dy[:live].double().sum(0)
            assert float((db.double() - wb).abs().max()) <= tol * float(dy[:live].double().abs().sum(0).max() if live else 0.0) + 1e-30
    dw2, dbs2, _ = call()
    assert torch.equal(dw, dw2) and all(torch.equal(a, c) for a, c in zip(dbs, dbs2))
    with pytest.raises(L.SgnnError):
        b = TE.WgradBatch()
        b.n_items, b.max_rows = 7, rows
        L.check(lib.sgnn_linear_wgrad_tc(C.byref(b), L.stream_ptr()))
