"""Host-side data path (sgnn_b200/data.py) against the reference's file format and sample indexing.
When the reference checkout is present (this container) its own loader is imported and compared
batch for batch; on a box without it the numpy restatement below is the check."""
import importlib.util
import os

import numpy as np
import pytest
import torch

from sgnn_b200 import data as D

REF_LOADER = "/root/reference/datasets/taylor_impact_2d/taylor_impact_data_loader.py"


def _write(tmp_path, n_traj=3, s=4):
    rng = np.random.default_rng(0)
    trs = {}
    for i in range(n_traj):
        t, n = 7 + i, 5 + 2 * i
        trs[f"case_{i}.npz"] = (rng.normal(size=(t, n, 2)).astype(np.float32),
                                np.full(n, i % 2, dtype=np.int64),
                                rng.normal(size=(t, n)).astype(np.float32))
    path = tmp_path / "train.npz"
    D.save_trajectories(str(path), trs, {"stress_mean": 1.5, "stress_std": 2.0})
    return path, list(trs.values())


def test_round_trip_and_stats(tmp_path):
    path, trs = _write(tmp_path)
    back = D.load_trajectories(str(path))
    assert len(back) == len(trs)
    for a, b in zip(back, trs):
        for x, y in zip(a, b):
            assert np.array_equal(x, y)
    assert D.stress_stats(str(path)) == {"mean": 1.5, "std": 2.0}
    md = D.compute_metadata(trs)
    assert md["dim"] == 2 and len(md["vel_mean"]) == 2 and len(md["acc_std"]) == 2
    with pytest.raises(FileNotFoundError):
        D.load_trajectories(str(tmp_path / "missing.npz"))


def test_sample_indexing_and_batch_layout(tmp_path):
    s = 4
    path, trs = _write(tmp_path, s=s)
    ds = D.DeviceSamples(str(path), s, device="cpu")
    assert len(ds) == sum(t[0].shape[0] - s for t in trs)
    idx = 0
    for ti, (pos, types, stress) in enumerate(trs):
        for t in range(s, pos.shape[0]):
            assert ds.locate(idx) == (ti, t)
            b = ds.batch([idx])
            assert np.array_equal(b["input"]["positions"].numpy(), np.transpose(pos[t - s:t], (1, 0, 2)))
            assert np.array_equal(b["output"]["next_position"].numpy(), pos[t])
            assert np.array_equal(b["output"]["next_strain"].numpy(), stress[t])
            assert b["input"]["particle_type"].tolist() == [int(types[0])] * pos.shape[1]
            idx += 1
    b = ds.batch([0, len(ds) - 1])
    n0, n1 = trs[0][0].shape[1], trs[-1][0].shape[1]
    assert b["input"]["n_particles_per_example"].tolist() == [n0, n1]
    assert b["input"]["positions"].shape == (n0 + n1, s, 2) and b["input"]["positions"].is_contiguous()
    assert b["meta"]["trajectory_idx"].tolist() == [0, len(trs) - 1]


@pytest.mark.skipif(not os.path.exists(REF_LOADER), reason="reference checkout not present")
def test_matches_reference_loader(tmp_path):
    spec = importlib.util.spec_from_file_location("ref_loader", REF_LOADER)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    s = 4
    path, _ = _write(tmp_path, s=s)
    rds = ref.TaylorImpactSamplesDataset(str(path), s)
    ds = D.DeviceSamples(str(path), s, device="cpu")
    assert len(rds) == len(ds)
    picks = [0, 3, len(ds) - 1, 5]
    rb = ref.collate_fn([rds[i] for i in picks])
    b = ds.batch(picks)
    for grp in ("input", "output", "meta"):
        assert set(rb[grp]) == set(b[grp])
        for k in rb[grp]:
            assert torch.equal(rb[grp][k].to(b[grp][k].dtype), b[grp][k]), (grp, k)
    rt = ref.TaylorImpactTrajectoriesDataset(str(path))
    tt = D.TrajectoriesOnDevice(str(path), device="cpu")
    assert len(rt) == len(tt)
    for i in range(len(tt)):
        a, c = rt[i], tt[i]
        assert set(a) == set(c)
        for k in ("positions", "particle_type", "n_particles_per_example", "strains"):
            assert torch.equal(a[k].to(c[k].dtype), c[k]), k


def test_loader_shards_cover_each_global_batch_once(tmp_path):
    path, _ = _write(tmp_path)
    ds = D.DeviceSamples(str(path), 4, device="cpu")
    world, bs = 2, 2
    seen = []
    loaders = [D.DeviceSampleLoader(ds, bs, True, 7, r, world) for r in range(world)]
    steps = [len(ld) for ld in loaders]
    assert steps[0] == steps[1] == len(ds) // (world * bs)
    for ld in loaders:
        for b in ld:
            seen += list(zip(b["meta"]["trajectory_idx"].tolist(), b["meta"]["time_idx"].tolist()))
    assert len(seen) == len(set(seen)) == steps[0] * world * bs
    one = D.DeviceSampleLoader(ds, 4, False)
    assert sum(b["input"]["n_particles_per_example"].numel() for b in one) == len(ds)


def test_synthetic_trajectory_frame0_on_lattice():
    pos, types, stress = D.synthetic_taylor_trajectory(6, 4, 0.5, frames=5)
    assert pos.shape == (5, 24, 2) and stress.shape == (5, 24) and types.shape == (24,)
    xs = np.unique(pos[0, :, 0])
    assert len(xs) == 6 and np.allclose(np.diff(xs), 0.5)
