"""GPU: spatially sharded rollout (sgnn_b200/sharding.py) vs the unsharded rollout on ONE GPU, the
ranks emulated as threads of this process (ThreadComm).  Edge sets of owned receivers must be
identical; positions agree to summation-order noise (tile boundaries fall differently)."""
import threading

import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("regime,world,precision", [("A", 3, "tf32"), ("B", 2, "fp32")])
def test_sharded_rollout_equals_unsharded(regime, world, precision):
    from sgnn_b200 import sharding as S, synthetic
    from sgnn_b200.single_scale import LearnedSimulator
    nx, ny, Sw, layers, nsteps = (60, 16, 6, 2, 5) if regime == "A" else (72, 12, 6, 2, 4)
    spacing, radius = synthetic.spacing_for(regime)
    traj = synthetic.taylor_trajectory(nx, ny, spacing, nframes=Sw + nsteps, seed=4)
    stats = synthetic.normalization_stats(synthetic.metadata_from(traj["positions"]), 0.02 * spacing / 0.5)
    def make():
        return LearnedSimulator(2, (Sw - 1) * 2 + 1, 3, 64, layers, 1, 64, radius, stats, 1, 9,
                                device="cuda", precision=precision).cuda().eval()
    torch.manual_seed(1)
    sim = make()
    pos = traj["positions"].cuda()
    n = pos.shape[0]
    # unsharded reference run
    cur = pos[:, :Sw].contiguous()
    want = []
    _, ei_full, _ = sim._encoder_preprocessor(cur, [n], traj["particle_types"].cuda())
    for _ in range(nsteps):
        nxt, _ = sim.predict_positions(cur, [n], traj["particle_types"].cuda())
        want.append(nxt)
        cur = torch.cat([cur[:, 1:], nxt[:, None]], 1)
    want = torch.stack(want)
    # sharded run, ranks as threads
    shared = S.ThreadComm.Shared(world)
    bounds = S.slab_bounds(n, world)
    sims = []
    for _ in range(world):                      # one simulator (own kernel plans/buffers) per rank
        sims.append(make())
        sims[-1].load_state_dict(sim.state_dict())
    results, errors = [None] * world, []

    def run(r):
        try:
            torch.cuda.set_device(0)
            lo, hi = bounds[r]
            results[r] = S.sharded_rollout(sims[r], pos[lo:hi, :Sw].contiguous(), nsteps,
                                           S.ThreadComm(shared, r), n_layers=layers, refresh_every=2,
                                           return_edges=True, use_graph=False,    # ranks are threads here
                                           fixed_ghosts=(regime == "A"))         # both ghost plans are covered
        except Exception as ex:      # surface failures instead of dead-locking the barrier
            errors.append(ex)
            shared.barrier.abort()

    th = [threading.Thread(target=run, args=(r,)) for r in range(world)]
    [t.start() for t in th]
    [t.join(timeout=120) for t in th]
    assert not errors, errors
    got = torch.cat([results[r][0] for r in range(world)], dim=1)
    edges = []
    for r in range(world):
        ei, nl = results[r][2]
        edges.append(ei - nl + bounds[r][0])
    assert torch.equal(torch.cat(edges, 1), ei_full), "edge sets of owned receivers differ"
    err = (got - want).abs().max().item()
    scale = pos.abs().max().item()
    print(f"sharded vs unsharded ({regime}, world {world}, {precision}): max |dpos| = {err:.3e}")
    assert err <= (2e-6 if precision == "fp32" else 2e-5) * scale


def test_graph_replayed_steps_equal_eager_steps():
    """The per-refresh CUDA graph over the fixed window buffer (what the multi-GPU run uses) against the
    eagerly launched loop and against the unsharded rollout, one rank."""
    from sgnn_b200 import sharding as S, synthetic
    from sgnn_b200.single_scale import LearnedSimulator
    nx, ny, Sw, layers, nsteps = 60, 16, 6, 2, 7
    spacing, radius = synthetic.spacing_for("A")
    traj = synthetic.taylor_trajectory(nx, ny, spacing, nframes=Sw + nsteps, seed=5)
    stats = synthetic.normalization_stats(synthetic.metadata_from(traj["positions"]), 0.02 * spacing / 0.5)
    torch.manual_seed(1)
    sim = LearnedSimulator(2, (Sw - 1) * 2 + 1, 3, 64, layers, 1, 64, radius, stats, 1, 9,
                           device="cuda", precision="tf32").cuda().eval()
    pos = traj["positions"].cuda()
    n = pos.shape[0]
    init = pos[:, :Sw].contiguous()
    comm = S.ThreadComm(S.ThreadComm.Shared(1), 0)
    eager, st_e = S.sharded_rollout(sim, init, nsteps, comm, n_layers=layers, refresh_every=3, use_graph=False)
    graph, st_g = S.sharded_rollout(sim, init, nsteps, comm, n_layers=layers, refresh_every=3, use_graph=True)
    assert torch.equal(eager, graph) and torch.equal(st_e, st_g)
    cur = init
    for k in range(nsteps):
        nxt, _ = sim.predict_positions(cur, [n], traj["particle_types"].cuda())
        assert torch.equal(nxt, graph[k]), k          # one rank, no ghosts: the very same launches
        cur = torch.cat([cur[:, 1:], nxt[:, None]], 1)
