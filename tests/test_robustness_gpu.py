"""GPU: lifetime and robustness cases (round-1 advisor findings).

  * a cached rollout graph must stay valid after the simulator's bounded engine cache evicted the
    engine it was captured on (plans own what they captured)
  * 64-wide tf32 model WITH a particle-type embedding (would otherwise pick the gather-mode kernels,
    which have no embedding prologue) vs the oracle
  * neighbour search on Inf / NaN / 1e30 positions: no fault, in-range output
"""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

S, H = 6, 64


def _traj(nx, ny, nsteps, seed=0, regime="A"):
    from sgnn_b200 import synthetic
    spacing, radius = synthetic.spacing_for(regime)
    traj = synthetic.taylor_trajectory(nx, ny, spacing, nframes=S + nsteps, seed=seed)
    stats = synthetic.normalization_stats(synthetic.metadata_from(traj["positions"]), 0.02 * spacing / 0.5)
    return traj, stats, radius


def test_cached_plan_survives_engine_eviction():
    from sgnn_b200.single_scale import LearnedSimulator, evaluate as EV
    EV.VERBOSE = False
    shapes = [(20, 12), (100, 60), (160, 75), (20, 12)]        # 240, 6000, 12000, 240 particles
    traj0, stats, radius = _traj(*shapes[0], nsteps=4)
    torch.manual_seed(1)
    sim = LearnedSimulator(2, (S - 1) * 2 + 1, 3, H, 3, 1, H, radius, stats, 1, 9, device="cuda").cuda().eval()
    outs = []
    for nx, ny in shapes:
        traj, _, _ = _traj(nx, ny, nsteps=4)
        pos = traj["positions"].cuda()
        out = EV.rollout(sim, pos, traj["particle_types"].cuda(), torch.tensor(pos.shape[0]),
                         traj["strains"].cuda(), 4, 2, "cuda", S)
        outs.append(out["predicted_rollout"])
    assert len(sim._engines) <= 2                                # the first engine was evicted ...
    assert np.array_equal(outs[0], outs[3])                      # ... and its plan still replays correctly
    assert np.isfinite(outs[3]).all()


@pytest.mark.parametrize("precision", ["tf32", "fp32"])
def test_type_embedding_h64_matches_oracle(precision):
    from oracle import sgnn_oracle as O
    from sgnn_b200.single_scale import LearnedSimulator, evaluate as EV
    EV.VERBOSE = False
    traj, stats, radius = _traj(24, 12, nsteps=3, regime="B")
    ntypes, emb = 2, 16
    torch.manual_seed(1)
    sim = LearnedSimulator(2, (S - 1) * 2 + 1 + emb, 3, H, 3, 1, H, radius, stats, ntypes, emb,
                           device="cuda", precision=precision).cuda().eval()
    pos = traj["positions"]
    n = pos.shape[0]
    types = (torch.arange(n) % 3 == 0).long()
    nxt, strain = sim.predict_positions(pos[:, :S].cuda(), [torch.tensor(n)], types.cuda())
    sd = {k: v.detach().cpu() for k, v in sim.state_dict().items()}
    want, want_strain, _ = O.predict_positions(O.cast_sd(sd, torch.float64), pos[:, :S].double(), [n], types, radius,
                                               O.cast_stats(stats, torch.float64))
    acc_std = stats["acceleration"]["std"].max().item()
    tol = {"fp32": 2e-5, "tf32": 3e-3}[precision]
    assert (nxt.cpu().double() - want).abs().max().item() <= tol * acc_std
    assert (strain.cpu().double() - want_strain).abs().max().item() <= tol
    # int32 / non-contiguous ids are converted, CPU ids are refused, the rollout driver takes the same path
    nxt2, _ = sim.predict_positions(pos[:, :S].cuda(), [torch.tensor(n)], types.int().cuda())
    assert torch.equal(nxt, nxt2)
    from sgnn_b200._lib import SgnnError
    with pytest.raises(SgnnError):
        sim.predict_positions(pos[:, :S].cuda(), [torch.tensor(n)], types)
    out = EV.rollout(sim, pos.cuda(), types.cuda(), torch.tensor(n), traj["strains"].cuda(), 3, 2, "cuda", S)
    assert np.array_equal(out["predicted_rollout"][0], nxt.cpu().numpy())


@pytest.mark.parametrize("bad", [float("inf"), float("-inf"), float("nan"), 1e30, -3e38])
@pytest.mark.parametrize("mode", [0, 1])
def test_radius_graph_survives_non_finite_positions(bad, mode):
    import ctypes as C
    from sgnn_b200 import _lib as L
    from sgnn_b200.engine import RadiusGraph
    from sgnn_b200.single_scale.learned_simulator import example_ptr_of
    from oracle.radius import radius_graph_np
    g = torch.Generator().manual_seed(0)
    pos = (torch.rand(2000, 2, generator=g) * 40).cuda()
    clean = pos.clone()
    pos[17, 0] = bad
    pos[900, 1] = bad
    old = L.lib().sgnn_radius_graph_mode(mode)
    try:
        rg = RadiusGraph(2000, 20, True, "cuda")
        csr = rg.build(pos, 2, 2, example_ptr_of([2000], "cuda"), 3.0)
        torch.cuda.synchronize()                                 # no fault
        e = int(csr.rowptr[2000])
        assert 0 <= e <= 2000 * 20
        if e:
            assert int(csr.src[:e].min()) >= 0 and int(csr.src[:e].max()) < 2000
        if np.isfinite(bad):                                     # huge but finite: still the exact edge set
            want = radius_graph_np(pos.cpu().numpy(), 3.0, [2000], loop=True, max_num_neighbors=20)
            assert torch.equal(csr.edge_index().cpu(), torch.from_numpy(want))
        # and the same buffers still give the exact answer on clean input afterwards
        csr = rg.build(clean, 2, 2, example_ptr_of([2000], "cuda"), 3.0)
        want = radius_graph_np(clean.cpu().numpy(), 3.0, [2000], loop=True, max_num_neighbors=20)
        assert torch.equal(csr.edge_index().cpu(), torch.from_numpy(want))
    finally:
        L.lib().sgnn_radius_graph_mode(old)
