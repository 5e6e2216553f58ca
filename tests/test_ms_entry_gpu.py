"""End-to-end through the multi-scale entry points (sgnn_b200/multi_scale/multi_scale_train.py,
multi_scale_inference.py, static_graph_data_loader.py): a tiny synthetic dataset on the 0.5-spaced lattice in
the reference's NPZ layout -> train() with validation and best-checkpoint saving -> resume -> run_inference in
rollout mode writing the reference's pickles -> the CLI (reference multi_scale_train.py:79-447,
multi_scale_inference.py:19-129)."""
import pickle
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

S = 4


def _dataset(root: Path):
    from sgnn_b200 import data as D
    shapes = [(12, 6), (10, 8), (12, 6)]
    trs = {f"case_{i}.npz": D.synthetic_taylor_trajectory(nx, ny, 0.5, frames=S + 5, seed=i)
           for i, (nx, ny) in enumerate(shapes)}
    md = D.compute_metadata(list(trs.values()), {"file_valid": ["case_2.npz"], "file_test": ["case_2.npz", "case_1.npz"]})
    D.save_trajectories(str(root / "train.npz"), {k: trs[k] for k in ("case_0.npz", "case_1.npz")}, md)
    D.save_trajectories(str(root / "valid.npz"), {"case_2.npz": trs["case_2.npz"]})
    D.save_trajectories(str(root / "test.npz"), {"case_2.npz": trs["case_2.npz"], "case_1.npz": trs["case_1.npz"]})
    return md


def _config(tmp_path, **kw):
    cfg = {"mode": "train", "data_path": str(tmp_path / "data"), "model_path": str(tmp_path / "models"),
           "output_path": str(tmp_path / "rollouts"), "num_scales": 2, "window_size": 2, "radius_multiplier": 2.0,
           "layers": 2, "hidden_dim": 128, "dim": 2, "particle_type_embedding_size": 9, "input_sequence_length": S,
           "batch_size": 2, "noise_std": 0.002, "ntraining_steps": 6, "nsave_steps": 3, "loss_weight_position": 1.0,
           "loss_weight_strain": 1.0, "lr_init": 1e-3, "lr_decay": 0.1, "lr_decay_steps": 100, "log": False,
           "project_name": "p", "run_name": "ms", "model_file": None, "train_state_file": None,
           "inference_mode": "autoregressive", "debug_graph": False}
    cfg.update(kw)
    return cfg


def test_static_graphs_match_the_oracle_and_are_cached_on_the_device(tmp_path):
    from oracle import sgnn_oracle_ms as OM
    from sgnn_b200 import data as D
    from sgnn_b200.multi_scale import static_graph_data_loader as SL
    _dataset(tmp_path / "data")
    trs = D.load_trajectories(str(tmp_path / "data" / "train.npz"))
    ds = SL.MultiScaleTaylorImpactSamplesDataset(str(tmp_path / "data" / "train.npz"), S, num_scales=2, window_size=2,
                                                 radius_multiplier=2.0)
    for i, tr in enumerate(trs):
        want = OM.create_all_edges(torch.from_numpy(np.asarray(tr[0][0], dtype=np.float32)), 2, 2, 2.0)
        for k in ("grid2mesh_edges", "mesh2mesh_edges", "mesh2grid_edges"):
            got = ds._static_graphs[i][k]
            assert got.is_cuda and torch.equal(got.cpu(), want[k]), (i, k)
    b = ds.batch([int(ds.starts[1]), 0])
    assert b["graph"] is ds._static_graphs[1] and b["input"]["positions"].is_cuda


def test_train_resume_inference(tmp_path):
    from sgnn_b200.multi_scale import multi_scale_train as MT
    md = _dataset(tmp_path / "data")
    cfg = _config(tmp_path)
    dev = torch.device("cuda")
    torch.manual_seed(0)
    sim = MT._get_simulator(md, cfg["noise_std"], cfg["noise_std"], dev, cfg)
    assert len([m for m in sim._multi_scale_gnn.g2m_block.edge_fn[0] if isinstance(m, torch.nn.Linear)]) == 3   # nmlp_layers = 2 (reference :331)
    before = {k: v.detach().cpu().clone() for k, v in sim.state_dict().items()}
    step, lowest = MT.train(sim, md, dev, cfg, log_every=3)
    assert step == 6 and np.isfinite(lowest)
    assert any(not torch.equal(before[k], v.detach().cpu()) for k, v in sim.state_dict().items())
    saved = sorted((tmp_path / "models" / "ms").glob("model-best-*.pt"))
    assert saved and saved[0].name == "model-best-000003.pt"
    st = torch.load(str(saved[-1]).replace("model-best", "train_state-best"))
    assert set(st) == {"optimizer_state", "global_train_state"}
    assert st["global_train_state"]["step"] == int(saved[-1].stem[-6:])

    cfg2 = _config(tmp_path, model_file=saved[-1].name,
                   train_state_file=saved[-1].name.replace("model-best", "train_state-best"),
                   ntraining_steps=int(saved[-1].stem[-6:]) + 2, nsave_steps=1000)
    sim2 = MT._get_simulator(md, cfg["noise_std"], cfg["noise_std"], dev, cfg2)
    step2, _ = MT.train(sim2, md, dev, cfg2)
    assert step2 == cfg2["ntraining_steps"]

    cfg3 = _config(tmp_path, mode="rollout", model_file=saved[-1].name)
    sim3 = MT._get_simulator(md, cfg["noise_std"], cfg["noise_std"], dev, cfg3)
    loss = MT.predict(sim3, md, dev, cfg3)
    assert np.isfinite(loss)
    shapes = {"case_2": (5, 72, 2), "case_1": (5, 80, 2)}
    for case, shape in shapes.items():        # two trajectories of different size: each rolled out on ITS graph
        with open(tmp_path / "rollouts" / "ms" / f"{case}.pkl", "rb") as f:
            out = pickle.load(f)
        for k in ("initial_positions", "predicted_rollout", "ground_truth_rollout", "predicted_strain", "particle_types",
                  "rmse_position", "rmse_strain", "metadata", "case_name", "inference_mode", "run_time"):
            assert k in out, k
        assert out["predicted_rollout"].shape == shape and np.isfinite(out["predicted_rollout"]).all()

    # a run without validation data keeps the final weights (reference :262-271)
    (tmp_path / "data" / "valid.npz").rename(tmp_path / "data" / "valid.off")
    cfg4 = _config(tmp_path, run_name="nov", ntraining_steps=2, nsave_steps=1)
    sim4 = MT._get_simulator(md, cfg["noise_std"], cfg["noise_std"], dev, cfg4)
    MT.train(sim4, md, dev, cfg4)
    assert (tmp_path / "models" / "nov" / "model-final-000002.pt").exists()
    assert (tmp_path / "models" / "nov" / "train_state-final-000002.pt").exists()


def test_cli_main(tmp_path):
    import yaml
    from sgnn_b200.multi_scale import multi_scale_train as MT
    _dataset(tmp_path / "data")
    cfg = _config(tmp_path, run_name="cli", layers=1, ntraining_steps=2, nsave_steps=2)
    p = tmp_path / "config.yaml"
    p.write_text(yaml.safe_dump(cfg))
    MT.main(["--config", str(p), "--mode", "train", "--log", "False"])
    assert (tmp_path / "models" / "cli" / "model-best-000002.pt").exists()
    metrics = MT.main(["--config", str(p), "--mode", "valid", "--model_file", "model-best-000002.pt"])
    assert set(metrics) == {"val/loss_total", "val/loss_position", "val/loss_strain", "val/loss_oneStep",
                            "val/mean_time", "val/std_time"}
    MT.main(["--config", str(p), "--mode", "rollout", "--model_file", "model-best-000002.pt"])
    assert (tmp_path / "rollouts" / "cli" / "case_2.pkl").exists()
