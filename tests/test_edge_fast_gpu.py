"""GPU: the throughput edge-layer and node kernels (4 tile workers per SM, hidden activations through TMEM)
must reproduce the one-tile-per-CTA kernel it replaces, on graphs with many tiles, ragged tails,
CSR rows straddling warps/tiles (degree 20) and nmlp_layers in {1, 2}."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu


def _sim(nmlp, radius, stats):
    from sgnn_b200.single_scale import LearnedSimulator
    torch.manual_seed(5)
    return LearnedSimulator(2, 21, 3, 64, 3, nmlp, 64, radius, stats, 1, 9, device="cuda").cuda().eval()


@pytest.mark.parametrize("regime,nx,ny", [("A", 150, 67), ("B", 96, 41), ("C", 40, 33)])
@pytest.mark.parametrize("nmlp", [1, 2])
def test_fast_equals_generic(regime, nx, ny, nmlp):
    from sgnn_b200 import synthetic
    spacing, radius = synthetic.spacing_for(regime)
    traj = synthetic.taylor_trajectory(nx, ny, spacing, nframes=12, seed=3)
    pos = traj["positions"].cuda()
    stats = synthetic.normalization_stats(synthetic.metadata_from(traj["positions"]), 0.02 * spacing / 0.5)
    os.environ["SGNN_EDGE_PQ"] = "1"       # this test is about the P/Q kernels; gather mode: test_edge_gather_gpu.py
    try:
        sim = _sim(nmlp, radius, stats)
        n = pos.shape[0]
        sim._engine(n, pos.device)         # the engine (and its mode) is built here
    finally:
        os.environ.pop("SGNN_EDGE_PQ", None)
    types = traj["particle_types"].cuda()
    os.environ.pop("SGNN_EDGE_GENERIC", None)
    os.environ.pop("SGNN_NODE_GENERIC", None)
    a, sa = sim.predict_positions(pos[:, :11], [n], types)
    b, sb = sim.predict_positions(pos[:, :11], [n], types)
    assert torch.equal(a, b)                                   # deterministic
    # The generic kernels with ONE thread per row (their default at 64-wide latents on the tensor cores) add the
    # LayerNorm moments in the order of the throughput kernels: the same bits.  With two threads per row the moments
    # are combined over the halves: a last-bit change before a TF32 rounding, i.e. TF32-level differences in the
    # outputs, bounded like any TF32 result (3e-3 acceleration standard deviations, tests/test_single_scale_gpu.py).
    acc_std = stats["acceleration"]["std"].max().item()
    for parts, tol_pos, tol_strain in ((None, 1e-6 * max(1.0, pos.abs().max().item()), 1e-4), ("2", 3e-3 * acc_std, 3e-3)):
        if parts:
            os.environ["SGNN_GENERIC_PARTS"] = parts
        try:
            a, sa = sim.predict_positions(pos[:, :11], [n], types)     # nmlp = 2: some of its layers run on the generic kernels too
            os.environ["SGNN_EDGE_GENERIC"] = "1"
            os.environ["SGNN_NODE_GENERIC"] = "1"
            os.environ["SGNN_ENC_GENERIC"] = "1"
            c, sc = sim.predict_positions(pos[:, :11], [n], types)
        finally:
            os.environ.pop("SGNN_EDGE_GENERIC", None)
            os.environ.pop("SGNN_NODE_GENERIC", None)
            os.environ.pop("SGNN_ENC_GENERIC", None)
            os.environ.pop("SGNN_GENERIC_PARTS", None)
        torch.cuda.synchronize()
        assert torch.isfinite(a).all()
        err = (a - c).abs().max().item()
        serr = (sa - sc).abs().max().item()
        print(f"regime {regime} nmlp {nmlp} parts {parts or 'default'}: N={n} max|fast-generic| pos {err:.3e} strain {serr:.3e}")
        assert err <= tol_pos and serr <= tol_strain, (parts, err, tol_pos, serr, tol_strain)
