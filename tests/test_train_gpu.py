"""GPU parity of the training path: predict_accelerations (noisy inputs, targets), the loss and
EVERY parameter gradient against the unmodified reference's fp64 run (golden fixtures), for the
single-scale and the multi-scale simulators.  The training path computes in fp32 (cuBLAS GEMMs with
TF32 off + the deterministic graph kernels), so the bar is the reference's own fp32-vs-fp64 gap
times a small factor: |g - g64| <= 20 * gap32 + 1e-6 * max|g64|."""
import pytest
import torch

from tests.util import Fixture

pytestmark = pytest.mark.gpu


def _check_grads(sim, f, loss_scale=1.0):
    g64 = f.grads()
    seen = 0
    for k, p in sim.named_parameters():
        if k not in g64:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, k
            continue
        gap, gmax = f[f"grad_gap/{k}"].tolist()
        err = (p.grad.cpu() * loss_scale - g64[k]).abs().max().item()
        assert err <= 20 * gap + 1e-6 * gmax + 1e-9, (k, err, gap, gmax)
        seen += 1
    assert seen == len(g64)


@pytest.mark.parametrize("name", ["single_h64", "single_h32_batch"])
def test_single_scale_train_step_matches_reference(name):
    from tests.test_single_scale_gpu import make_sim
    f = Fixture(name)
    sim, S, radius, _ = make_sim(f, "tf32")
    sim.train()
    pos, noise = f.t("train/positions").cuda(), f.t("train/noise").cuda()
    nxt, nstrain = f.t("train/next_position").cuda(), f.t("train/next_strain").cuda()
    counts, types = f.t("train/counts").cuda(), f.t("train/particle_types").cuda()
    pa, ta, ps = sim.predict_accelerations(nxt, noise, pos, counts, types)
    assert torch.allclose(ta.cpu(), f.t("train/target_acc_f32"), atol=1e-4, rtol=1e-4)
    assert (pa.cpu().double() - f.t("train/pred_acc_f64")).abs().max() <= 5e-5
    assert (ps.cpu().double() - f.t("train/pred_strain_f64")).abs().max() <= 5e-5
    loss = (((pa - ta) ** 2).sum(-1) + (ps - nstrain) ** 2).mean()
    assert abs(loss.item() - float(f["train/loss_f64"])) <= 1e-4 * max(1.0, float(f["train/loss_f64"]))
    sim.zero_grad()
    loss.backward()
    _check_grads(sim, f)
    # deterministic backward (no atomics)
    g1 = [p.grad.clone() for p in sim.parameters() if p.grad is not None]
    sim.zero_grad()
    pa, ta, ps = sim.predict_accelerations(nxt, noise, pos, counts, types)
    (((pa - ta) ** 2).sum(-1) + (ps - nstrain) ** 2).mean().backward()
    g2 = [p.grad for p in sim.parameters() if p.grad is not None]
    assert all(torch.equal(a, b) for a, b in zip(g1, g2))


@pytest.mark.parametrize("name", ["multi_ref_fixture", "multi_h128"])
def test_multi_scale_train_step_matches_reference(name):
    from tests.test_multi_scale_gpu import graph_of, make_sim
    f = Fixture(name)
    sim, S, _, _ = make_sim(f, "tf32")
    sim.set_static_graph(graph_of(f))
    sim.train()
    pos, types = f.t("positions").cuda(), f.t("particle_types").cuda()
    n = pos.shape[0]
    noise = f.t("train/noise").cuda()
    pa, ta, ps = sim.predict_accelerations(pos[:, S], noise, pos[:, :S], torch.tensor([n]), types)
    assert torch.allclose(ta.cpu(), f.t("train/target_acc_f32"), atol=1e-4, rtol=1e-4)
    assert (pa.cpu().double() - f.t("train/pred_acc_f64")).abs().max() <= 5e-5
    loss = (((pa - ta) ** 2).sum(-1) + ps ** 2).mean()
    assert abs(loss.item() - float(f["train/loss_f64"])) <= 1e-4 * max(1.0, float(f["train/loss_f64"]))
    sim.zero_grad()
    loss.backward()
    _check_grads(sim, f)


def test_train_step_api_decreases_loss_and_graph_ops_gradcheck():
    """training.train_step end to end (noise on device, Adam), plus a finite-difference check of
    the two autograd Functions in fp32 against a dense torch restatement."""
    from sgnn_b200 import train_ops as T, training
    from tests.test_single_scale_gpu import make_sim
    f = Fixture("single_h64")
    sim, S, radius, _ = make_sim(f, "tf32")
    sim.train()
    batch = {"positions": f.t("train/positions").cuda(), "particle_type": f.t("train/particle_types").cuda(),
             "n_particles_per_example": f.t("train/counts").cuda(),
             "next_position": f.t("train/next_position").cuda(), "next_strain": f.t("train/next_strain").cuda()}
    opt = torch.optim.Adam(sim.parameters(), lr=1e-3)
    noise = f.t("train/noise").cuda()
    losses = [training.train_step(sim, opt, batch, 0.02, noise=noise).item() for _ in range(8)]
    assert losses[-1] < losses[0]
    assert abs(losses[0] - float(f["train/loss_f64"])) <= 1e-4 * max(1.0, float(f["train/loss_f64"]))
    # graph ops vs dense torch
    g = torch.Generator().manual_seed(0)
    n, e, w = 50, 400, 32
    tgt = torch.sort(torch.randint(0, n, (e,), generator=g)).values
    src = torch.randint(0, n, (e,), generator=g)
    rowptr = torch.zeros(n + 1, dtype=torch.int32)
    rowptr[1:] = torch.cumsum(torch.bincount(tgt, minlength=n), 0).int()
    graph = T.TrainGraph(rowptr.cuda(), src.int().cuda(), tgt.int().cuda(), n, e)
    p = torch.randn(n, w, generator=g).cuda().requires_grad_()
    q = torch.randn(n, w, generator=g).cuda().requires_grad_()
    wgt = torch.randn(e, w, generator=g).cuda()
    out = T.segment_sum(T.gather_add(p, q, graph) * wgt, graph)
    out.square().sum().backward()
    p2, q2 = p.detach().clone().requires_grad_(), q.detach().clone().requires_grad_()
    ref = torch.zeros(n, w, device="cuda").index_add_(0, tgt.cuda(), (p2[tgt.cuda()] + q2[src.cuda()]) * wgt)
    ref.square().sum().backward()
    assert torch.allclose(out, ref, atol=1e-4, rtol=1e-5)
    assert torch.allclose(p.grad, p2.grad, atol=1e-3, rtol=1e-4)
    assert torch.allclose(q.grad, q2.grad, atol=1e-3, rtol=1e-4)


@pytest.mark.parametrize("width", [32, 64, 128])
@pytest.mark.parametrize("rows", [1, 77, 20011])
def test_layer_norm_kernels_match_aten(width, rows):
    """csrc/train_norm.cu forward / backward against F.layer_norm in fp64, plus bit-reproducibility."""
    import torch.nn.functional as F
    from sgnn_b200 import train_ops as T
    g = torch.Generator().manual_seed(width + rows)
    ln = torch.nn.LayerNorm(width).cuda()
    with torch.no_grad():
        ln.weight.copy_(torch.randn(width, generator=g))
        ln.bias.copy_(torch.randn(width, generator=g))
    x = (3.0 * torch.randn(rows, width, generator=g) + 1.5).cuda().requires_grad_()
    up = torch.randn(rows, width, generator=g).cuda()
    y = T.layer_norm(ln, x)
    (y * up).sum().backward()
    got = [y.detach(), x.grad.clone(), ln.weight.grad.clone(), ln.bias.grad.clone()]
    x64 = x.detach().double().requires_grad_()
    w64, b64 = ln.weight.detach().double().requires_grad_(), ln.bias.detach().double().requires_grad_()
    y64 = F.layer_norm(x64, (width,), w64, b64, ln.eps)
    (y64 * up.double()).sum().backward()
    want = [y64.detach(), x64.grad, w64.grad, b64.grad]
    for a, b, name in zip(got, want, ["y", "dx", "dgamma", "dbeta"]):
        tol = 2e-5 * max(1.0, float(b.abs().max()))
        assert (a.double() - b).abs().max().item() <= tol, name
    ln.zero_grad()
    x.grad = None
    (T.layer_norm(ln, x) * up).sum().backward()
    assert torch.equal(x.grad, got[1]) and torch.equal(ln.weight.grad, got[2]) and torch.equal(ln.bias.grad, got[3])


@pytest.mark.parametrize("rows,m,k", [(0, 64, 64), (1, 64, 64), (33, 64, 64), (20011, 64, 64), (4097, 64, 30),
                                      (3001, 64, 3), (2500, 2, 64), (1500, 128, 128), (777, 96, 130), (1000, 7, 5),
                                      (150000, 64, 64)])
@pytest.mark.parametrize("bias", [True, False])
def test_linear_weight_gradient_kernel(rows, m, k, bias):
    """csrc/train_linear.cu: dW = dy^T x and db = column sums of dy against fp64, input gradient against
    F.linear's, weight slices (non-contiguous views) as the interaction blocks pass them, bit-reproducible."""
    import torch.nn.functional as F
    from sgnn_b200 import train_ops as T
    g = torch.Generator().manual_seed(rows + 7 * m + k)
    full = torch.randn(m, k + 5, generator=g).cuda().requires_grad_()
    b = torch.randn(m, generator=g).cuda().requires_grad_() if bias else None
    x = (torch.randn(rows, k, generator=g) + 0.25).cuda().requires_grad_()
    up = torch.randn(rows, m, generator=g).cuda()

    def run():
        full.grad = x.grad = None
        if b is not None:
            b.grad = None
        y = T.linear(x, full[:, 2:2 + k], b)
        (y * up).sum().backward()
        return y.detach(), x.grad.clone(), full.grad.clone(), (b.grad.clone() if b is not None else None)

    y, dx, dw, db = run()
    assert torch.equal(y, F.linear(x.detach(), full.detach()[:, 2:2 + k], b.detach() if b is not None else None))
    w64 = full.detach().double()[:, 2:2 + k]
    want_dw = up.double().t() @ x.detach().double()
    scale = max(1.0, float(want_dw.abs().max())) if rows else 1.0
    assert (dw[:, 2:2 + k].double() - want_dw).abs().max().item() <= 2e-6 * scale + 1e-5
    assert dw[:, :2].abs().max().item() == 0 and dw[:, 2 + k:].abs().max().item() == 0
    assert torch.allclose(dx.double(), up.double() @ w64, rtol=1e-5, atol=1e-4)
    if bias:
        want_db = up.double().sum(0)
        assert (db.double() - want_db).abs().max().item() <= 1e-5 * max(1.0, rows ** 0.5)
    again = run()
    assert torch.equal(again[2], dw) and (db is None or torch.equal(again[3], db))


@pytest.mark.parametrize("n,deg,ni,ne,h,bias", [(500, 4, 64, 64, 64, True), (301, 7, 32, 64, 128, True),
                                                (64, 1, 64, 32, 64, False)])
def test_fused_edge_first_linear_matches_concatenation(n, deg, ni, ne, h, bias):
    """train_ops.edge_pre against Linear(cat([x_i, x_j, e])) in fp64 (forward and all four gradients)."""
    from sgnn_b200 import train_ops as T
    from sgnn_b200.engine import Csr
    g = torch.Generator().manual_seed(n + deg)
    tgt = torch.randint(0, n, (n * deg,), generator=g)
    src = torch.randint(0, n, (n * deg,), generator=g)
    csr, _ = Csr.from_edge_index(torch.stack([src, tgt]).cuda(), n)
    graph = T.TrainGraph.from_csr(csr)
    x = torch.randn(n, ni, generator=g).cuda().requires_grad_()
    e = torch.randn(graph.e, ne, generator=g).cuda().requires_grad_()
    w = (0.2 * torch.randn(h, 2 * ni + ne, generator=g)).cuda().requires_grad_()
    b = torch.randn(h, generator=g).cuda().requires_grad_() if bias else None
    up = torch.randn(graph.e, h, generator=g).cuda()
    pre = T.edge_pre(x, e, w, b, graph)
    (pre * up).sum().backward()
    x64, e64, w64 = (t.detach().double().requires_grad_() for t in (x, e, w))
    b64 = b.detach().double().requires_grad_() if bias else None
    cat = torch.cat([x64[graph.tgt.long()], x64[graph.src.long()], e64], 1)
    want = torch.nn.functional.linear(cat, w64, b64)
    (want * up.double()).sum().backward()
    pairs = [(pre.detach(), want.detach(), "pre"), (x.grad, x64.grad, "dx"), (e.grad, e64.grad, "de"),
             (w.grad, w64.grad, "dw")] + ([(b.grad, b64.grad, "db")] if bias else [])
    for a, c, name in pairs:
        tol = 2e-5 * max(1.0, float(c.abs().max()))
        assert (a.double() - c).abs().max().item() <= tol, name


def test_tf32_matmul_mode_of_train_step_stays_close():
    """training.train_step(matmul="tf32"): same loss to 1e-3 relative, the global flag is restored."""
    from sgnn_b200 import training
    from tests.test_single_scale_gpu import make_sim
    f = Fixture("single_h64")
    batch = {"positions": f.t("train/positions").cuda(), "particle_type": f.t("train/particle_types").cuda(),
             "n_particles_per_example": f.t("train/counts").cuda(),
             "next_position": f.t("train/next_position").cuda(), "next_strain": f.t("train/next_strain").cuda()}
    noise = f.t("train/noise").cuda()
    losses = {}
    for mode in ("fp32", "tf32"):
        sim, _, _, _ = make_sim(f, "tf32")
        sim.train()
        opt = torch.optim.SGD(sim.parameters(), lr=0.0)
        before = torch.backends.cuda.matmul.allow_tf32
        losses[mode] = training.train_step(sim, opt, batch, 0.02, noise=noise, matmul=mode).item()
        assert torch.backends.cuda.matmul.allow_tf32 == before
    assert abs(losses["tf32"] - losses["fp32"]) <= 1e-3 * max(1.0, abs(losses["fp32"]))


def test_two_forwards_before_one_backward_keep_their_graphs():
    """predict_accelerations hands each forward its own copy of the edge list: a second forward (gradient
    accumulation, an extra loss evaluation) must not change the gradients of the first."""
    from sgnn_b200 import synthetic
    from sgnn_b200.single_scale import LearnedSimulator
    s = 6
    spacing, radius = synthetic.spacing_for("B")
    stats = None
    batches = []
    for seed in (0, 1):
        traj = synthetic.taylor_trajectory(20, 10, spacing, nframes=s + 1, seed=seed)
        stats = stats or synthetic.normalization_stats(synthetic.metadata_from(traj["positions"]), 0.02 * spacing / 0.5)
        pos = traj["positions"].cuda()
        if seed == 1:
            pos = pos * 0.8                                   # a different edge set
        batches.append((pos[:, :s].contiguous(), pos[:, s].contiguous()))
    torch.manual_seed(0)
    sim = LearnedSimulator(2, (s - 1) * 2 + 1, 3, 64, 2, 1, 64, radius, stats, 1, 9, device="cuda").cuda().train()
    n = batches[0][0].shape[0]
    types, counts = torch.zeros(n, dtype=torch.int64, device="cuda"), torch.tensor([n], device="cuda")

    def fwd(b):
        pa, ta, ps = sim.predict_accelerations(b[1], torch.zeros_like(b[0]), b[0], counts, types)
        return ((pa - ta) ** 2).sum()

    sim.zero_grad()
    fwd(batches[0]).backward()
    want = {k: p.grad.clone() for k, p in sim.named_parameters() if p.grad is not None}
    sim.zero_grad()
    l0 = fwd(batches[0])
    fwd(batches[1])                                           # overwrites the persistent neighbour-search buffers
    l0.backward()
    for k, p in sim.named_parameters():
        if p.grad is not None:
            assert torch.equal(p.grad, want[k]), k
