"""Gather-mode edge layer (edge_layer_ws.cu: sender rows by TMA gather4, K = 128 first GEMM,
warp-specialised pipeline) against the P/Q formulation (edge_layer_fast.cu) on the same weights and
inputs, and against the CPU oracle.  The two device paths differ only in where the sender term is
rounded to TF32 (x before the product vs. the fp32 product Q), so they must agree far inside the
tolerance the TF32 path has against the fp32 oracle.  Sizes cover: fewer tiles than SMs, a ragged
last tile, and several tiles per CTA and per epilogue group (ring wrap-around, barrier phases)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

S, H, LAYERS = 11, 64, 5


def _case(nx, ny, regime, seed=0):
    from sgnn_b200 import synthetic
    spacing, radius = synthetic.spacing_for(regime)
    traj = synthetic.taylor_trajectory(nx, ny, spacing, nframes=S + 2, seed=seed)
    stats = synthetic.normalization_stats(synthetic.metadata_from(traj["positions"]), 0.02 * spacing / 0.5)
    return traj, stats, radius


def _sim(stats, radius, state=None):
    from sgnn_b200.single_scale import LearnedSimulator
    torch.manual_seed(1)
    sim = LearnedSimulator(2, (S - 1) * 2 + 1, 3, H, LAYERS, 1, H, radius, stats, 1, 9,
                           device="cuda", precision="tf32").cuda().eval()
    if state is not None:
        sim.load_state_dict(state)
    return sim


@pytest.mark.parametrize("nx,ny,regime", [(24, 12, "A"), (60, 40, "B"), (200, 160, "A"), (120, 80, "B")])
def test_gather_mode_matches_pq_mode(monkeypatch, nx, ny, regime):
    traj, stats, radius = _case(nx, ny, regime)
    pos = traj["positions"][:, :S].cuda()
    types = traj["particle_types"].cuda()
    n = pos.shape[0]
    sim_g = _sim(stats, radius)
    monkeypatch.setenv("SGNN_KEEP_X", "1")       # gather mode keeps fp32 latents on request only
    nxt_g, strain_g = sim_g.predict_positions(pos, [torch.tensor(n)], types)
    eng, _ = sim_g._engine(n, pos.device)
    assert eng.gather, "gather mode was expected for the 64-wide single-scale network"
    x_g = eng.x_final.clone()
    monkeypatch.delenv("SGNN_KEEP_X")
    monkeypatch.setenv("SGNN_EDGE_PQ", "1")
    sim_p = _sim(stats, radius, sim_g.state_dict())
    nxt_p, strain_p = sim_p.predict_positions(pos, [torch.tensor(n)], types)
    eng_p, _ = sim_p._engine(n, pos.device)
    assert not eng_p.gather
    x_p = eng_p.x_final
    torch.cuda.synchronize()
    # node latents after 5 layers are O(1..10); TF32 has 2^-11 relative rounding per operand
    scale = x_p.abs().max().item()
    assert (x_g - x_p).abs().max().item() <= 4e-3 * scale, ((x_g - x_p).abs().max().item(), scale)
    acc_std = stats["acceleration"]["std"].max().item()
    assert (nxt_g - nxt_p).abs().max().item() <= 1e-2 * acc_std + 1e-6
    assert (strain_g - strain_p).abs().max().item() <= 1e-2


def test_gather_mode_matches_oracle():
    from oracle import sgnn_oracle as O
    traj, stats, radius = _case(40, 24, "B")
    pos, types = traj["positions"][:, :S], traj["particle_types"]
    n = pos.shape[0]
    sim = _sim(stats, radius)
    nxt, strain = sim.predict_positions(pos.cuda(), [torch.tensor(n)], types.cuda())
    assert sim._engine(n, torch.device("cuda"))[0].gather
    sd = {k: v.detach().cpu() for k, v in sim.state_dict().items()}
    want, want_strain, _ = O.predict_positions(sd, pos, [n], types, radius, stats)
    acc_std = stats["acceleration"]["std"].max().item()
    # stated one-step tolerance of the tf32 path: 3e-3 of acc_std + the fp32 quantum of the position
    assert (nxt.cpu() - want).abs().max().item() <= 3e-3 * acc_std + float(np.spacing(np.float32(pos.abs().max().item())))
    assert (strain.cpu() - want_strain).abs().max().item() <= 3e-3


def test_gather_mode_rollout_is_deterministic():
    """Two 5-step rollouts give bit-identical trajectories (fixed tile -> group assignment, ordered
    segmented sums: no atomics anywhere)."""
    from sgnn_b200.single_scale import evaluate as EV
    traj, stats, radius = _case(60, 40, "A", seed=3)
    from sgnn_b200 import synthetic
    spacing, _ = synthetic.spacing_for("A")
    full = synthetic.taylor_trajectory(60, 40, spacing, nframes=S + 5, seed=3)
    sim = _sim(stats, radius)
    n = full["positions"].shape[0]
    outs = []
    for _ in range(2):
        o = EV.rollout(sim, full["positions"].cuda(), full["particle_types"].cuda(), torch.tensor(n),
                       full["strains"].cuda(), 5, 2, torch.device("cuda"), S)
        outs.append(o["predicted_rollout"].copy())
    assert (outs[0] == outs[1]).all()
