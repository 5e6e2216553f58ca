"""The hand-written training step (sgnn_b200/train_engine.py, csrc/train_mlp.cu) against (a) the autograd path
of train_ops.py on the same inputs, (b) the oracle's autograd in fp64 on BASELINE configs[2] (8 x 4800
particles; reference sgnn/single_scale/train.py:230-280), (c) itself: CUDA-graph replay == eager launches."""
import copy

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

S, H = 6, 64


def _batch(samples, nx, ny, regime="B", seed0=0):
    from sgnn_b200 import synthetic
    spacing, radius = synthetic.spacing_for(regime)
    pos, nxt, strain = [], [], []
    for k in range(samples):
        traj = synthetic.taylor_trajectory(nx, ny, spacing, nframes=S + 1, seed=seed0 + k)
        pos.append(traj["positions"][:, :S]); nxt.append(traj["positions"][:, S]); strain.append(traj["strains"][S])
    stats = synthetic.normalization_stats(synthetic.metadata_from(traj["positions"]), 0.02 * spacing / 0.5)
    n = pos[0].shape[0]
    batch = {"positions": torch.cat(pos).cuda(), "next_position": torch.cat(nxt).cuda(),
             "next_strain": torch.cat(strain).cuda(), "particle_type": torch.zeros(n * samples, dtype=torch.int64).cuda(),
             "n_particles_per_example": torch.full((samples,), n, dtype=torch.int64).cuda()}
    return batch, stats, radius, spacing


def _sim(stats, radius, layers):
    from sgnn_b200.single_scale import LearnedSimulator
    torch.manual_seed(3)
    return LearnedSimulator(2, (S - 1) * 2 + 1, 3, H, layers, 1, H, radius, stats, 1, 9, device="cuda").cuda().train()


def _autograd_grads(sim, batch, noise):
    from sgnn_b200 import training
    for p in sim.parameters():
        p.grad = None
    pa, ta, ps = sim.predict_accelerations(batch["next_position"], noise, batch["positions"],
                                           batch["n_particles_per_example"], batch["particle_type"])
    loss = training.loss_terms(pa, ta, ps, batch["next_strain"]).sum() / batch["positions"].shape[0]
    loss.backward()
    return loss.item(), {k: p.grad.detach().clone() for k, p in sim.named_parameters() if p.grad is not None}


@pytest.mark.parametrize("precision,tol", [("fp32", 2e-5), ("tf32", 3e-2)])
def test_engine_gradients_match_autograd_path(precision, tol):
    from sgnn_b200 import noise_utils, train_engine, training
    batch, stats, radius, spacing = _batch(2, 24, 12)
    sim = _sim(stats, radius, 2)
    assert train_engine.supports(sim)
    noise = noise_utils.get_random_walk_noise_for_position_sequence(batch["positions"], 0.02 * spacing,
                                                                    torch.Generator(device="cuda").manual_seed(5))
    want_loss, want = _autograd_grads(sim, batch, noise)
    opt = training.make_optimizer(sim.parameters(), 1e-3)
    eng = train_engine.TrainEngine(sim, opt, batch["positions"].shape[0], "cuda", precision=precision, use_graph=False)
    loss = eng.forward_backward(batch, noise).item()
    assert abs(loss - want_loss) <= (1e-5 if precision == "fp32" else 5e-3) * max(1.0, want_loss), (loss, want_loss)
    worst = 0.0
    for k, p in sim.named_parameters():
        if k not in want:
            assert float(p.grad.abs().max()) == 0.0, k
            continue
        err = (p.grad - want[k]).abs().max().item()
        scale = want[k].abs().max().item()
        worst = max(worst, err / (tol * scale + 1e-12))
        assert err <= tol * scale + 1e-9, (k, err, scale)
    again = eng.forward_backward(batch, noise).item()          # deterministic kernels: same bits
    assert again == loss
    print(f"\n{precision}: loss {loss:.6f} (autograd {want_loss:.6f}); worst gradient error / bar = {worst:.3f}")


def test_engine_config3_gradients_vs_fp64_oracle():
    """BASELINE configs[2], one step on 8 x 4800 particles: every gradient within 20 x the reference arithmetic's
    own fp32-vs-fp64 gap (the bar of tests/test_baseline_configs_gpu.py::test_config3_train_step_8x4800)."""
    from oracle import sgnn_oracle as O
    from sgnn_b200 import train_engine, training
    from tests import test_baseline_configs_gpu as B
    samples = 8
    pos, nxt, strain = [], [], []
    for k in range(samples):
        traj, stats, radius = B._case("A", 1, seed=100 + k)
        pos.append(traj["positions"][:, :B.S]); nxt.append(traj["positions"][:, B.S]); strain.append(traj["strains"][B.S])
    pos, nxt, strain = torch.cat(pos), torch.cat(nxt), torch.cat(strain)
    n = pos.shape[0]
    counts = torch.full((samples,), n // samples, dtype=torch.int64)
    types = torch.zeros(n, dtype=torch.int64)
    noise = O.random_walk_noise(pos, 0.02 * 25.0, torch.Generator().manual_seed(2))
    sim = B._sim(stats, radius, "tf32").train()
    batch = {"positions": pos.cuda(), "next_position": nxt.cuda(), "next_strain": strain.cuda(),
             "particle_type": types.cuda(), "n_particles_per_example": counts.cuda()}
    eng = train_engine.TrainEngine(sim, training.make_optimizer(sim.parameters(), 1e-4), n, "cuda", use_graph=False)
    loss = eng.forward_backward(batch, noise.cuda()).item()
    sd = {k: v.detach().cpu() for k, v in sim.state_dict().items()}

    def oracle_grads(dtype):
        sdd = {k: v.clone().requires_grad_(v.is_floating_point()) for k, v in O.cast_sd(sd, dtype).items()}
        opa, ota, ops, ei = O.predict_accelerations(sdd, nxt.to(dtype), noise.to(dtype), pos.to(dtype), counts, types,
                                                    radius, O.cast_stats(stats, dtype))
        ol = O.train_loss(opa, ota, ops, strain.to(dtype))
        ol.backward()
        return ol.item(), {k: v.grad for k, v in sdd.items() if v.requires_grad and v.grad is not None}

    l64, g64 = oracle_grads(torch.float64)
    _, g32 = oracle_grads(torch.float32)
    assert abs(loss - l64) <= 1e-4 * max(1.0, l64), (loss, l64)
    worst, seen = 0.0, 0
    for k, p in sim.named_parameters():
        if k not in g64:
            assert float(p.grad.abs().max()) == 0.0, k
            continue
        gap = (g32[k].double() - g64[k]).abs().max().item()
        gmax = g64[k].abs().max().item()
        err = (p.grad.cpu().double() - g64[k]).abs().max().item()
        worst = max(worst, err / (20 * gap + 1e-6 * gmax + 1e-12))
        assert err <= 20 * gap + 1e-6 * gmax + 1e-12, (k, err, gap, gmax)
        seen += 1
    assert seen == len(g64)
    print(f"\nconfig 3 (engine, hi/lo-split TF32): N={n} loss {loss:.6f} (fp64 {l64:.6f}); "
          f"worst gradient error / bar = {worst:.3f} over {seen} tensors")


def test_engine_graph_replay_equals_eager_and_tracks_autograd():
    from sgnn_b200 import train_engine, training
    batch, stats, radius, spacing = _batch(2, 24, 12, seed0=7)
    n = batch["positions"].shape[0]
    sims = [_sim(stats, radius, 2) for _ in range(3)]
    for s in sims[1:]:
        s.load_state_dict(sims[0].state_dict())
    opts = [training.make_optimizer(s.parameters(), 1e-3) for s in sims]
    eng_graph = train_engine.TrainEngine(sims[0], opts[0], n, "cuda", use_graph=True)
    eng_eager = train_engine.TrainEngine(sims[1], opts[1], n, "cuda", use_graph=False)
    losses = [[], [], []]
    for step in range(4):
        gens = [torch.Generator(device="cuda").manual_seed(100 + step) for _ in range(3)]
        losses[0].append(eng_graph.step(batch, 0.02 * spacing, generator=gens[0]).item())
        losses[1].append(eng_eager.step(batch, 0.02 * spacing, generator=gens[1]).item())
        losses[2].append(training.train_step(sims[2], opts[2], batch, 0.02 * spacing, generator=gens[2]).item())
    assert losses[0] == losses[1], (losses[0], losses[1])
    for (k, a), (_, b) in zip(sims[0].state_dict().items(), sims[1].state_dict().items()):
        assert torch.equal(a, b), k
    assert np.allclose(losses[0], losses[2], rtol=2e-4), (losses[0], losses[2])
    for (k, a), (_, c) in zip(sims[0].state_dict().items(), sims[2].state_dict().items()):
        if a.is_floating_point():    # Adam normalises the gradient: elements with |g| ~ 0 may move by lr in either direction
            assert (a - c).abs().mean().item() <= 1e-4, k
    assert losses[0][-1] < losses[0][0]
    assert float(opts[0].state[next(iter(sims[0].parameters()))]["step"]) == 4.0
