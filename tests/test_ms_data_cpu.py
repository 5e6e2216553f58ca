"""Host-side logic of the multi-scale data path (sgnn_b200/multi_scale/static_graph_data_loader.py) against
the reference's item / batch layout (static_graph_data_loader.py:63-232).  The graph builder needs the GPU, so
here it is swapped for the oracle's CPU restatement (tests may use the oracle; the product never does)."""
import importlib.util
import os

import numpy as np
import pytest
import torch

from sgnn_b200 import data as D
from sgnn_b200.multi_scale import static_graph_data_loader as SL

REF_LOADER = "/root/reference/datasets/taylor_impact_2d/taylor_impact_data_loader.py"
KEYS = ("grid2mesh_edges", "mesh2mesh_edges", "mesh2grid_edges")


@pytest.fixture
def cpu_graphs(monkeypatch):
    from oracle import sgnn_oracle_ms as OM
    monkeypatch.setattr(SL, "build_static_multi_scale_graph",
                        lambda p0, num_scales=3, window_size=3, radius_multiplier=2.0:
                        OM.create_all_edges(p0.cpu(), num_scales, window_size, radius_multiplier))
    return OM


def _write(tmp_path, shapes=((8, 4), (6, 6), (10, 4)), frames=8):
    trs = {f"case_{i}.npz": D.synthetic_taylor_trajectory(nx, ny, 0.5, frames=frames, seed=i)
           for i, (nx, ny) in enumerate(shapes)}
    path = tmp_path / "train.npz"
    D.save_trajectories(str(path), trs, {"stress_mean": 0.0, "stress_std": 1.0})
    return path, list(trs.values())


def test_samples_carry_their_trajectory_graph(tmp_path, cpu_graphs):
    s = 3
    path, trs = _write(tmp_path)
    ds = SL.MultiScaleTaylorImpactSamplesDataset(str(path), s, num_scales=2, window_size=2, radius_multiplier=2.0,
                                                 device="cpu")
    assert len(ds) == sum(t[0].shape[0] - s for t in trs)
    want = [cpu_graphs.create_all_edges(torch.from_numpy(np.asarray(t[0][0], dtype=np.float32)), 2, 2, 2.0) for t in trs]
    for idx in (0, 4, 5, 9, 10, len(ds) - 1):
        tr, t = ds.locate(idx)
        item = ds[idx]
        assert item["meta"] == {"trajectory_idx": tr, "time_idx": t}
        assert np.array_equal(item["input"]["positions"].numpy(), np.transpose(trs[tr][0][t - s:t], (1, 0, 2)))
        assert np.array_equal(item["output"]["next_position"].numpy(), trs[tr][0][t])
        assert item["input"]["n_particles_per_example"] == trs[tr][0].shape[1]
        for k in KEYS:
            assert torch.equal(item["graph"][k], want[tr][k]), k
        assert set(item["graph"]["graph_hierarchy"]) == {0, 1}


def test_collate_takes_the_first_samples_graph(tmp_path, cpu_graphs):
    """Reference quirk (static_graph_data_loader.py:226-229): whatever trajectories a batch mixes, it carries
    the graph of batch[0]."""
    s = 3
    path, trs = _write(tmp_path)
    ds = SL.MultiScaleTaylorImpactSamplesDataset(str(path), s, num_scales=2, window_size=2, device="cpu")
    first_of = [int(ds.starts[i]) for i in range(len(trs))]
    for order in ([first_of[1], first_of[0]], [first_of[2], first_of[1], first_of[0]]):
        b = ds.batch(order)
        assert b["graph"] is ds._static_graphs[ds.locate(order[0])[0]]
        c = SL.multi_scale_collate_fn([ds[i] for i in order])
        assert c["graph"] is b["graph"]
        for grp in ("input", "output", "meta"):
            for k in b[grp]:
                assert torch.equal(b[grp][k].cpu(), c[grp][k].cpu()), (grp, k)
        assert b["input"]["n_particles_per_example"].tolist() == [trs[ds.locate(i)[0]][0].shape[1] for i in order]


@pytest.mark.skipif(not os.path.exists(REF_LOADER), reason="reference checkout not present")
def test_batches_match_the_reference_collate(tmp_path, cpu_graphs):
    spec = importlib.util.spec_from_file_location("ref_loader", REF_LOADER)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    s = 3
    path, _ = _write(tmp_path)
    ours = SL.MultiScaleTaylorImpactSamplesDataset(str(path), s, num_scales=2, window_size=2, device="cpu")
    theirs = ref.TaylorImpactSamplesDataset(str(path), s)
    assert len(ours) == len(theirs)
    for order in ([0, 7, 12], [len(ours) - 1, 3]):
        a = ours.batch(order)
        b = ref.collate_fn([theirs[i] for i in order])
        for grp in ("input", "output"):
            for k in b[grp]:
                assert np.array_equal(np.asarray(a[grp][k]), np.asarray(b[grp][k])), (grp, k)


def test_loader_shards_by_rank_and_trajectories_iterate(tmp_path, cpu_graphs):
    s = 3
    path, trs = _write(tmp_path)
    seen = []
    for rank in range(2):
        ld = SL.get_multi_scale_data_loader_by_samples(str(path), s, batch_size=2, shuffle=True, num_scales=2,
                                                       window_size=2, device="cpu", seed=5, rank=rank, world=2)
        batches = list(ld)
        assert len(batches) == len(ld) == len(ld.samples) // 4
        assert all("graph" in b for b in batches)
        seen.append([tuple(zip(b["meta"]["trajectory_idx"].tolist(), b["meta"]["time_idx"].tolist())) for b in batches])
    flat0, flat1 = [x for b in seen[0] for x in b], [x for b in seen[1] for x in b]
    assert not set(flat0) & set(flat1)
    trajs = SL.get_multi_scale_data_loader_by_trajectories(str(path), num_scales=2, window_size=2, device="cpu")
    items = list(trajs)
    assert len(items) == len(trs)
    for i, it in enumerate(items):
        assert it["positions"].shape == (trs[i][0].shape[1], trs[i][0].shape[0], 2)
        assert it["strains"].shape == trs[i][2].shape and it["trajectory_idx"] == i
        assert it["graph"] is trajs._static_graphs[i]


def test_config_file_has_the_reference_keys():
    from sgnn_b200.multi_scale import multi_scale_train as MT
    cfg = MT.load_config("multi_scale_config.yaml")          # resolved beside the module, like the reference
    want = {"mode", "data_path", "model_path", "output_path", "num_scales", "window_size", "radius_multiplier", "layers",
            "hidden_dim", "dim", "particle_type_embedding_size", "input_sequence_length", "batch_size", "noise_std",
            "ntraining_steps", "nsave_steps", "loss_weight_position", "loss_weight_strain", "lr_init", "lr_decay",
            "lr_decay_steps", "log", "project_name", "run_name", "model_file", "train_state_file", "inference_mode",
            "debug_graph"}
    assert set(cfg) == want
    with pytest.raises(FileNotFoundError):
        MT.load_config("no_such_config.yaml")
