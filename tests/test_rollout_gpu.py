"""GPU: device-resident rollout driver (CUDA-graph replay) vs golden fixtures / oracle."""
import numpy as np
import pytest
import torch

from tests.test_single_scale_gpu import TOL, make_sim
from tests.util import Fixture

pytestmark = pytest.mark.gpu


def _run(sim, f, S, mode="autoregressive", graph=True, types=None):
    from sgnn_b200.single_scale import evaluate as EV
    EV.VERBOSE = False
    pos, strains = f.t("positions").cuda(), f.t("strains").cuda()
    types = f.t("particle_types").cuda() if types is None else types
    return EV.rollout(sim, pos, types, torch.tensor(pos.shape[0]), strains, 0, 2, "cuda", S, mode,
                      use_cuda_graph=graph)


@pytest.mark.parametrize("precision", ["fp32", "tf32"])
def test_rollout_golden_and_graph_determinism(precision):
    f = Fixture("single_h64")
    sim, S, radius, nsteps = make_sim(f, precision)
    a = _run(sim, f, S, graph=True)
    b = _run(sim, f, S, graph=True)          # second call: pure graph replay
    c = _run(sim, f, S, graph=False)
    assert np.array_equal(a["predicted_rollout"], b["predicted_rollout"])
    assert np.array_equal(a["predicted_rollout"], c["predicted_rollout"])   # deterministic kernels
    assert np.array_equal(a["predicted_strain"], c["predicted_strain"])
    ref64 = f["rollout_f64"]
    acc_std = f.stats()["acceleration"]["std"].max().item()
    assert a["predicted_rollout"].shape == ref64.shape
    from tests.test_single_scale_gpu import growth_bounds
    ulp = float(np.spacing(np.float32(np.abs(f["positions"]).max())))
    bounds, _ = growth_bounds(f["rollout_f32"], ref64, (5e-6 if precision == "fp32" else 3e-3) * acc_std + ulp)
    for step in range(nsteps if precision == "fp32" else 1):     # tf32: the trajectory leaves the fixture's at the first
        err = np.abs(a["predicted_rollout"][step] - ref64[step]).max()   # edge flip (tests/test_single_scale_gpu.py checks its
        assert err <= bounds[step], (step, err, bounds[step])            # steps teacher-forced against the oracle)
    if precision == "fp32":
        assert np.allclose(a["rmse_position"], f["rmse_position_f32"], rtol=1e-3)
    for k in ("initial_positions", "initial_strains", "ground_truth_rollout", "ground_truth_strain",
              "predicted_strain", "particle_types", "rmse_strain", "run_time", "inference_mode"):
        assert k in a
    assert a["initial_positions"].shape == (S, 240, 2)


def test_one_step_mode_matches_oracle():
    from oracle import sgnn_oracle as O
    f = Fixture("single_h64")
    sim, S, radius, nsteps = make_sim(f, "fp32")
    out = _run(sim, f, S, mode="one_step")
    pred, _, _ = O.rollout(f.state_dict(), f.t("positions"), f.t("particle_types"), nsteps, radius,
                           f.stats(), S, mode="one_step")
    assert np.abs(out["predicted_rollout"] - pred.numpy()).max() <= 2e-4
    with pytest.raises(ValueError):
        _run(sim, f, S, mode="bogus")


def test_erosional_particles_follow_ground_truth():
    f = Fixture("single_h64")
    sim, S, radius, nsteps = make_sim(f, "fp32")
    types = f.t("particle_types").clone()
    types[::7] = -1
    out = _run(sim, f, S, types=types.cuda())
    gt = f["positions"][:, S:].transpose(1, 0, 2)
    assert np.array_equal(out["predicted_rollout"][:, ::7], gt[:, ::7])
    assert not np.array_equal(out["predicted_rollout"][:, 1::7], gt[:, 1::7])


def test_input_validation():
    from sgnn_b200.single_scale import evaluate as EV
    f = Fixture("single_h64")
    sim, S, _, _ = make_sim(f, "tf32")
    pos, st, ty = f.t("positions").cuda(), f.t("strains").cuda(), f.t("particle_types").cuda()
    with pytest.raises(ValueError):
        EV.rollout(sim, pos[:, 0], ty, torch.tensor(240), st, 0, 2, "cuda", S)
    with pytest.raises(ValueError):
        EV.rollout(sim, pos, ty, torch.tensor(240), st[:, :10], 0, 2, "cuda", S)
    with pytest.raises(ValueError):
        EV.rollout(sim, pos[:, :5], ty, torch.tensor(240), st, 0, 2, "cuda", S)


def test_batched_rollout_equals_separate_rollouts():
    """A [B] tensor of particle counts rolls B trajectories out as one system (one CUDA graph): examples never
    interact, so each one matches its own rollout (segment sums are cut at different tile boundaries: not bitwise)."""
    from sgnn_b200 import synthetic
    from sgnn_b200.single_scale import LearnedSimulator, evaluate as EV
    EV.VERBOSE = False
    s, nsteps = 6, 8
    spacing, radius = synthetic.spacing_for("A")
    trajs = [synthetic.taylor_trajectory(30, 10 + 2 * k, spacing, nframes=s + nsteps, seed=20 + k) for k in range(3)]
    stats = synthetic.normalization_stats(synthetic.metadata_from(trajs[0]["positions"]), 0.02 * spacing / 0.5)
    torch.manual_seed(1)
    sim = LearnedSimulator(2, (s - 1) * 2 + 1, 3, 64, 3, 1, 64, radius, stats, 1, 9, device="cuda").cuda().eval()
    counts = torch.tensor([t["positions"].shape[0] for t in trajs], device="cuda")
    pos = torch.cat([t["positions"] for t in trajs]).cuda()
    strains = torch.cat([t["strains"] for t in trajs], dim=1).cuda()
    types = torch.zeros(pos.shape[0], dtype=torch.int64, device="cuda")
    both = EV.rollout(sim, pos, types, counts, strains, nsteps, 2, "cuda", s)["predicted_rollout"]
    off = 0
    for t in trajs:
        n = t["positions"].shape[0]
        one = EV.rollout(sim, t["positions"].cuda(), types[:n], torch.tensor(n), t["strains"].cuda(), nsteps, 2, "cuda", s)
        err = np.abs(both[:, off:off + n] - one["predicted_rollout"]).max()
        assert err <= 2e-3 * spacing, err
        off += n
