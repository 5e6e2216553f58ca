"""CPU tests: the oracle (oracle/) against the golden fixtures generated from the unmodified
reference (oracle/make_golden.py) and against the reference's own golden vectors
(sgnn/multi_scale/test/test_multi_scale_graph.py:101-140,325-348)."""
import numpy as np
import pytest
import torch

from oracle import sgnn_oracle as O
from oracle import sgnn_oracle_ms as OM
from oracle.radius import radius_graph_np, radius_graph_py
from tests.util import Fixture

torch.set_num_threads(1)


@pytest.mark.parametrize("case", ["lattice_exact", "lattice_trunc", "random", "ragged_batch",
                                  "dim3", "single", "duplicates"])
def test_radius_cases(case):
    f = Fixture("radius_cases")
    pos, counts, r = f[case + "/pos"], f[case + "/counts"].tolist(), float(f[case + "/r"][0])
    want = f[case + "/edge_index"]
    for method in ("brute", "cells"):
        got = radius_graph_np(pos, r, counts, True, 20, method=method)
        assert np.array_equal(got, want), method
    if pos.shape[0] <= 250:
        assert np.array_equal(radius_graph_py(pos, r, counts, True, 20), want)
    # target-sorted, sources ascending within a target, no cross-example edge
    assert np.all(np.diff(want[1]) >= 0)
    ptr = np.concatenate([[0], np.cumsum(counts)])
    ex = np.searchsorted(ptr, np.arange(pos.shape[0]), side="right") - 1
    assert np.array_equal(ex[want[0]], ex[want[1]])


def test_radius_strict_boundary():
    # 0.5-spaced lattice, r = 1.5: pairs at exactly d = 1.5 (3 lattice steps) are excluded
    f = Fixture("radius_cases")
    pos, ei = f["lattice_exact/pos"], f["lattice_exact/edge_index"]
    d2 = ((pos[ei[0]] - pos[ei[1]]) ** 2).sum(-1)
    assert d2.max() < 1.5 ** 2
    assert np.isclose(d2.max(), 2.0)  # (1,1) lattice steps = sqrt(2); (1.5,0) is out


def test_radius_loop_false_rule():
    rng = np.random.default_rng(3)
    pos = rng.uniform(0, 5, (200, 2)).astype(np.float32)
    a = radius_graph_np(pos, 1.0, None, loop=False, max_num_neighbors=4, method="brute")
    b = radius_graph_np(pos, 1.0, None, loop=False, max_num_neighbors=4, method="cells")
    c = radius_graph_py(pos, 1.0, None, loop=False, max_num_neighbors=4)
    assert np.array_equal(a, b) and np.array_equal(a, c)
    assert not np.any(a[0] == a[1])


@pytest.mark.parametrize("name", ["single_h64", "single_h32_batch"])
def test_single_scale_matches_reference(name):
    f = Fixture(name)
    nx, ny, S, H, L, nmlp, nsteps, ntypes, emb = f["meta"].tolist()
    _, radius = f["spacing_radius"].tolist()
    sd, stats = f.state_dict(), f.stats()
    pos, types = f.t("positions"), f.t("particle_types")
    ei0 = O.graph_connectivity(pos[:, S - 1], [pos.shape[0]], radius)
    assert torch.equal(ei0, f.t("edge_index/0"))
    nf, ef = O.single_scale_features(pos[:, :S], ei0, radius, stats, types,
                                     sd["_particle_type_embedding.weight"])
    assert torch.equal(nf, f.t("node_features_f32"))
    assert torch.equal(ef, f.t("edge_features_f32"))
    out, lat, e_last = O.encode_process_decode(sd, "_encode_process_decode.", nf, ei0, ef, True)
    assert torch.allclose(out, f.t("epd_out_f32"), atol=1e-5, rtol=1e-5)
    # the edge-latent doubling quirk (graph_network.py:173-176): e_L = 2^L * e_0 exactly
    e0 = O.mlp_ln(sd, "_encode_process_decode._encoder.edge_fn.", ef)
    assert torch.equal(e_last, e0 * (2.0 ** L))
    pred, strain, eis = O.rollout(sd, pos, types, nsteps, radius, stats, S)
    for s in range(nsteps):
        assert torch.equal(eis[s], f.t(f"edge_index/{s}")), s
    assert torch.allclose(pred, f.t("rollout_f32"), atol=1e-4, rtol=1e-5)
    assert torch.allclose(strain, f.t("rollout_strain_f32"), atol=1e-4, rtol=1e-4)
    # fp64 oracle vs fp64 reference: tight
    sd64, st64 = O.cast_sd(sd, torch.float64), O.cast_stats(stats, torch.float64)
    pred64, _, _ = O.rollout(sd64, pos.double(), types, nsteps, radius, st64, S)
    assert torch.allclose(pred64, f.t("rollout_f64"), atol=1e-9, rtol=1e-9)
    assert np.allclose(O.rollout_rmse(pred.numpy(), pos[:, S:].permute(1, 0, 2).numpy()),
                       f["rmse_position_f32"], rtol=1e-4)


@pytest.mark.parametrize("name", ["single_h64", "single_h32_batch"])
def test_single_scale_train_step_matches_reference(name):
    f = Fixture(name)
    _, radius = f["spacing_radius"].tolist()
    sd = {k: v.double().requires_grad_(v.is_floating_point()) for k, v in f.state_dict().items()}
    st64 = O.cast_stats(f.stats(), torch.float64)
    pa, ta, ps, _ = O.predict_accelerations(
        sd, f.t("train/next_position").double(), f.t("train/noise").double(),
        f.t("train/positions").double(), f["train/counts"].tolist(), f.t("train/particle_types"),
        radius, st64)
    assert torch.allclose(pa, f.t("train/pred_acc_f64"), atol=1e-9)
    assert torch.allclose(ta, f.t("train/target_acc_f64"), atol=1e-9)
    loss = O.train_loss(pa, ta, ps, f.t("train/next_strain").double())
    assert abs(loss.item() - float(f["train/loss_f64"])) < 1e-9
    loss.backward()
    for k, g in f.grads().items():
        assert torch.allclose(sd[k].grad.float(), g, atol=1e-6, rtol=1e-4), k


def test_noise_matches_reference():
    f = Fixture("single_h64")
    spacing, _ = f["spacing_radius"].tolist()
    torch.manual_seed(2)
    noise = O.random_walk_noise(f.t("train/positions"), 0.02 * spacing / 0.5)
    assert torch.equal(noise, f.t("train/noise"))


def test_hierarchy_golden_vectors_of_reference_tests():
    # test_multi_scale_graph.py:75-79,101-140: 4x4 unit lattice, 3 scales, window 2
    x = torch.arange(4, dtype=torch.float32)
    xx, yy = torch.meshgrid(x, x, indexing="ij")
    pos = torch.stack([xx.flatten(), yy.flatten()], 1)
    h = OM.build_hierarchy(pos, 3, 2)
    assert h[0]["sampling_indices"].tolist() == list(range(16))
    assert h[1]["sampling_indices"].tolist() == [0, 2, 8, 10]
    assert h[2]["sampling_indices"].tolist() == [0]
    assert [h[s]["spacing"] for s in range(3)] == [0.5, 1.0, 2.0]
    # :325-348: 16x16, 5 scales -> 256/64/16/4/1
    x = torch.arange(16, dtype=torch.float32)
    xx, yy = torch.meshgrid(x, x, indexing="ij")
    h = OM.build_hierarchy(torch.stack([xx.flatten(), yy.flatten()], 1), 5, 2)
    assert [h[s]["num_particles"] for s in range(5)] == [256, 64, 16, 4, 1]
    with pytest.raises(ValueError):
        OM.build_hierarchy(pos, 1, 2)


@pytest.mark.parametrize("name", ["multi_ref_fixture", "multi_h128"])
def test_multi_scale_matches_reference(name):
    f = Fixture(name)
    S, H, nedge_out, L, nmlp, num_scales, window, ntypes, emb, nsteps = f["meta"].tolist()
    mult = float(f["mult"][0])
    sd, stats = f.state_dict(), f.stats()
    pos, types = f.t("positions"), f.t("particle_types")
    g = OM.create_all_edges(pos[:, 0], num_scales, window, mult)
    assert torch.equal(g["grid2mesh_edges"], f.t("g2m"))
    assert torch.equal(g["mesh2mesh_edges"], f.t("m2m"))
    assert torch.equal(g["mesh2grid_edges"], f.t("m2g"))
    for s in range(num_scales):
        assert torch.equal(g["graph_hierarchy"][s]["sampling_indices"],
                           f.t(f"hier/{s}/sampling_indices"))
    nf, ef = OM.multi_scale_features(pos[:, :S], g, stats, num_scales, mult, types,
                                     sd["_particle_type_embedding.weight"])
    assert torch.equal(nf, f.t("node_features_f32"))
    for k in ("g2m", "m2m", "m2g"):
        assert torch.equal(ef[k], f.t(f"edge_features_f32/{k}"))
    pred, strain = OM.rollout(sd, pos, types, nsteps, g, stats, S, num_scales, mult)
    assert torch.allclose(pred, f.t("rollout_f32"), atol=1e-5, rtol=1e-5)
    assert torch.allclose(strain, f.t("rollout_strain_f32"), atol=1e-4, rtol=1e-4)
    sd64 = {k: v.double().requires_grad_(v.is_floating_point()) for k, v in sd.items()}
    st64 = O.cast_stats(stats, torch.float64)
    pa, ta, ps = OM.predict_accelerations(sd64, pos[:, S].double(), f.t("train/noise").double(),
                                          pos[:, :S].double(), types, g, st64, num_scales, mult)
    assert torch.allclose(pa, f.t("train/pred_acc_f64"), atol=1e-9)
    loss = O.train_loss(pa, ta, ps, torch.zeros_like(ps))
    loss.backward()
    for k, gr in f.grads().items():
        assert torch.allclose(sd64[k].grad.float(), gr, atol=1e-6, rtol=1e-4), k
