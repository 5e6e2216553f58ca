"""GPU: the opt-in single-cluster neighbour search (sgnn_radius_graph_mode(1); n <= 8192: one launch,
phases separated by the hardware cluster barrier, hand-offs through distributed shared memory) against
the oracle -- edge sets must be bit-exact, as on the default nine-launch path."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def cluster_mode():
    from sgnn_b200 import _lib as L
    old = L.lib().sgnn_radius_graph_mode(1)
    yield
    L.lib().sgnn_radius_graph_mode(old)


def _edges(rg, pos, counts, r):
    from sgnn_b200.single_scale.learned_simulator import example_ptr_of
    csr = rg.build(pos, pos.shape[1], pos.shape[1], example_ptr_of(list(counts), pos.device), r)
    torch.cuda.synchronize()
    return csr.edge_index().cpu().numpy()


def test_one_launch_below_8193_particles_and_repeated_calls():
    from oracle.radius import radius_graph_np
    from sgnn_b200 import _lib as L
    from sgnn_b200.engine import RadiusGraph
    assert L.lib().sgnn_radius_graph_launches(4800, 20) == 1
    assert L.lib().sgnn_radius_graph_launches(8192, 20) == 1
    assert L.lib().sgnn_radius_graph_launches(8193, 20) > 1
    rng = np.random.default_rng(5)
    n = 4800
    rg = RadiusGraph(n, 20, True, torch.device("cuda"))
    for it in range(6):          # different clouds (and grids) through the same buffers
        scale = (5.0, 30.0, 80.0)[it % 3]
        pos = rng.uniform(-scale, scale, (n, 2)).astype(np.float32)
        got = _edges(rg, torch.as_tensor(pos).cuda(), [n], 1.1)
        assert np.array_equal(got, radius_graph_np(pos, 1.1, [n], True, 20)), it


@pytest.mark.parametrize("n,dim,counts,k,loop", [(1, 2, [1], 20, True), (127, 2, [127], 20, True),
                                                  (513, 1, [513], 20, True), (129, 3, [100, 29], 24, False),
                                                  (8192, 2, [8192], 20, True), (7000, 3, [3000, 1, 3999], 8, True),
                                                  (2000, 2, [2000], 20, True)])
def test_shapes_vs_oracle(n, dim, counts, k, loop):
    from oracle.radius import radius_graph_np
    from sgnn_b200.engine import RadiusGraph
    rng = np.random.default_rng(n + dim)
    side = max(2.0, (n / 2.0) ** (1.0 / dim))
    if n == 2000:
        side = 4.0                      # dense: every receiver truncated to K, hundreds of candidates per cell row
    pos = rng.uniform(0, side, (n, dim)).astype(np.float32)
    rg = RadiusGraph(n, k, loop, torch.device("cuda"))
    got = _edges(rg, torch.as_tensor(pos).cuda(), counts, 1.3)
    assert np.array_equal(got, radius_graph_np(pos, 1.3, counts, loop, k))


def test_default_is_the_general_path():
    from oracle.radius import radius_graph_np
    from sgnn_b200 import _lib as L
    from sgnn_b200.engine import RadiusGraph
    L.lib().sgnn_radius_graph_mode(0)
    assert L.lib().sgnn_radius_graph_launches(4800, 20) > 1
    pos = np.random.default_rng(1).uniform(0, 40, (4800, 2)).astype(np.float32)
    rg = RadiusGraph(4800, 20, True, torch.device("cuda"))
    got = _edges(rg, torch.as_tensor(pos).cuda(), [4800], 1.2)
    assert np.array_equal(got, radius_graph_np(pos, 1.2, [4800], True, 20))
