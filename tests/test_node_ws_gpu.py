"""Warp-specialised node-layer kernel (node_ws.cu: [agg | x] operand by TMA tile loads, one issuer warp
per GEMM, four epilogue groups) against the 4-worker kernel (node_fast.cu) on the same weights and
inputs.  Both run the same GEMMs in the same order on the same TF32-rounded operands, so the results
must be BIT-identical; the env switch SGNN_NODE_FAST=1 selects the older kernel per launch.  Sizes
cover fewer tiles than SMs, a ragged last tile, several tiles per CTA and per epilogue group (ring
wrap-around, barrier phases); regime B graphs have CSR rows that straddle edge tiles at almost every
tile boundary, and the isolated-particle case has many receivers whose only edge is the self loop."""
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


def _sim(stats, radius):
    from sgnn_b200.single_scale import LearnedSimulator
    torch.manual_seed(1)
    return LearnedSimulator(2, (S - 1) * 2 + 1, 3, H, LAYERS, 1, H, radius, stats, 1, 9,
                            device="cuda", precision="tf32").cuda().eval()


def _run(sim, pos, types):
    n = pos.shape[0]
    nxt, strain = sim.predict_positions(pos, [torch.tensor(n)], types)
    eng, _ = sim._engine(n, pos.device)
    torch.cuda.synchronize()
    x = eng.x_final.clone() if eng.x_final is not None else None
    return nxt.clone(), strain.clone(), x, eng.pq.clone(), eng


@pytest.mark.parametrize("nx,ny,regime", [(24, 12, "A"), (60, 40, "B"), (200, 160, "A"), (400, 300, "B")])
def test_node_ws_is_bit_identical_to_node_fast(monkeypatch, nx, ny, regime):
    traj, stats, radius = _case(nx, ny, regime)
    pos = traj["positions"][:, :S].cuda()
    types = traj["particle_types"].cuda()
    sim = _sim(stats, radius)
    # product configuration: the latents travel only as the rounded copy in pq, no fp32 x buffers
    nxt_0, strain_0, x_0, pq_0, eng = _run(sim, pos, types)
    assert eng.gather, "gather mode (and with it node_ws_kernel) was expected for the 64-wide network"
    assert x_0 is None
    monkeypatch.setenv("SGNN_KEEP_X", "1")       # also write the fp32 latents, to compare them
    nxt_w, strain_w, x_w, pq_w, _ = _run(sim, pos, types)
    assert torch.equal(nxt_0, nxt_w) and torch.equal(strain_0, strain_w) and torch.equal(pq_0, pq_w)
    monkeypatch.delenv("SGNN_KEEP_X")
    monkeypatch.setenv("SGNN_NODE_FAST", "1")
    nxt_f, strain_f, x_f, pq_f, _ = _run(sim, pos, types)
    assert torch.equal(x_w, x_f)
    assert torch.equal(pq_w, pq_f)
    assert torch.equal(nxt_w, nxt_f)
    assert torch.equal(strain_w, strain_f)
    assert torch.isfinite(nxt_w).all()


def test_node_ws_isolated_particles(monkeypatch):
    """Every seventh particle is moved onto a loose lattice (3 radii apart) beside the block: its CSR
    row holds the self loop only, so many 128-edge tiles hold dozens of one-edge receivers."""
    traj, stats, radius = _case(48, 32, "A")
    pos = traj["positions"][:, :S].clone()
    k = torch.arange(pos[::7].shape[0], dtype=torch.float32)
    corner = pos[:, -1].max(dim=0).values + 5.0 * radius
    lattice = torch.stack([corner[0] + (k % 16) * 3.0 * radius, corner[1] + (k // 16) * 3.0 * radius], dim=1)
    pos[::7] = lattice[:, None, :] + (pos[::7] - pos[::7, -1:])    # keep each window's velocities
    pos, types = pos.cuda(), traj["particle_types"].cuda()
    sim = _sim(stats, radius)
    monkeypatch.setenv("SGNN_KEEP_X", "1")
    nxt_w, strain_w, x_w, _, eng = _run(sim, pos, types)
    assert eng.gather
    monkeypatch.delenv("SGNN_KEEP_X")
    monkeypatch.setenv("SGNN_NODE_FAST", "1")
    nxt_f, strain_f, x_f, _, _ = _run(sim, pos, types)
    assert torch.equal(x_w, x_f) and torch.equal(nxt_w, nxt_f) and torch.equal(strain_w, strain_f)
    assert torch.isfinite(x_w).all()
