"""GPU: neighbour search through the C ABI vs the oracle -- edge sets must be bit-exact."""
import numpy as np
import pytest
import torch

from tests.util import Fixture

pytestmark = pytest.mark.gpu


def gpu_edges(pos, counts, r, k=20, loop=True):
    from sgnn_b200.engine import RadiusGraph
    from sgnn_b200.single_scale.learned_simulator import example_ptr_of
    p = torch.as_tensor(pos, dtype=torch.float32).cuda().contiguous()
    rg = RadiusGraph(p.shape[0], k, loop, p.device)
    csr = rg.build(p, p.shape[1], p.shape[1], example_ptr_of(list(counts), p.device), r)
    torch.cuda.synchronize()
    return csr.edge_index().cpu().numpy(), csr


@pytest.mark.parametrize("case", ["lattice_exact", "lattice_trunc", "random", "ragged_batch",
                                  "dim3", "single", "duplicates"])
def test_golden_cases(case):
    f = Fixture("radius_cases")
    got, _ = gpu_edges(f[case + "/pos"], f[case + "/counts"].tolist(), float(f[case + "/r"][0]))
    assert np.array_equal(got, f[case + "/edge_index"])


@pytest.mark.parametrize("n,r,k,loop", [(5000, 0.9, 20, True), (5000, 2.5, 24, True),
                                         (3000, 1.3, 8, False), (20000, 0.35, 20, True)])
def test_random_vs_oracle(n, r, k, loop):
    from oracle.radius import radius_graph_np
    rng = np.random.default_rng(n + k)
    pos = rng.uniform(0, 30, (n, 2)).astype(np.float32)
    got, _ = gpu_edges(pos, [n], r, k, loop)
    want = radius_graph_np(pos, r, [n], loop, k)
    assert np.array_equal(got, want)


def test_taylor_configs_vs_oracle():
    from oracle.radius import radius_graph_np
    from sgnn_b200 import synthetic
    for cfg in ("A", "B", "C", "yaml"):
        s, r = synthetic.spacing_for(cfg)
        traj = synthetic.taylor_trajectory(120, 40, s, nframes=3, seed=1)
        pos = traj["positions"][:, 2].numpy()
        got, _ = gpu_edges(pos, [pos.shape[0]], r)
        want = radius_graph_np(pos, r, [pos.shape[0]], True, 20)
        assert np.array_equal(got, want), cfg


def test_batch_of_8_and_determinism():
    from oracle.radius import radius_graph_np
    from sgnn_b200 import synthetic
    traj = synthetic.taylor_trajectory(60, 20, 12.5, nframes=9, seed=3)
    pos = torch.cat([traj["positions"][:, t] for t in range(1, 9)]).numpy()
    counts = [1200] * 8
    a, _ = gpu_edges(pos, counts, 15.0)
    b, _ = gpu_edges(pos, counts, 15.0)
    assert np.array_equal(a, b)
    assert np.array_equal(a, radius_graph_np(pos, 15.0, counts, True, 20))


def test_million_particles_properties():
    """Full-size (config 5) check by size-independent properties + oracle on a sample."""
    from oracle.radius import radius_graph_np
    from sgnn_b200 import synthetic
    s, r = synthetic.spacing_for("A")
    pos = synthetic.taylor_bar_lattice(2048, 512, s)
    g = torch.Generator().manual_seed(0)
    pos = (pos + 0.01 * s * torch.randn(pos.shape, generator=g)).numpy()
    got, csr = gpu_edges(pos, [pos.shape[0]], r)
    src, tgt = got
    assert np.all(np.diff(tgt) >= 0)                                   # receiver-sorted
    same = np.diff(tgt) == 0
    assert np.all(np.diff(src)[same] > 0)                              # sources ascending, unique
    d2 = ((pos[src] - pos[tgt]) ** 2).sum(-1)
    assert d2.max() < r * r * (1 + 1e-6)
    assert np.bincount(tgt, minlength=pos.shape[0]).max() <= 20
    assert np.all(np.isin(np.arange(pos.shape[0]), src[src == tgt]))   # every self loop kept
    want = radius_graph_np(pos, r, [pos.shape[0]], True, 20, method="cells")
    assert np.array_equal(got, want)
