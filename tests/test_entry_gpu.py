"""End-to-end through the entry points (sgnn_b200/single_scale/train.py): a tiny synthetic dataset in
the reference's NPZ layout -> train() with periodic validation and best-checkpoint saving -> resume
-> predict() in rollout mode writing the reference's pickle."""
import pickle
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _dataset(root: Path, s: int):
    from sgnn_b200 import data as D
    trs = {f"case_{i}.npz": D.synthetic_taylor_trajectory(12, 6, 12.5, frames=s + 6, seed=i) for i in range(3)}
    md = D.compute_metadata(list(trs.values()), {"file_valid": ["case_2.npz"], "file_test": ["case_2.npz"]})
    D.save_trajectories(str(root / "train.npz"), {k: trs[k] for k in ("case_0.npz", "case_1.npz")}, md)
    D.save_trajectories(str(root / "valid.npz"), {"case_2.npz": trs["case_2.npz"]})
    D.save_trajectories(str(root / "test.npz"), {"case_2.npz": trs["case_2.npz"]})
    return md


def test_train_resume_predict(tmp_path):
    from sgnn_b200.single_scale import train as T
    s = 6
    md = _dataset(tmp_path / "data", s)
    cfg = {"mode": "train", "data_path": str(tmp_path / "data"), "model_path": str(tmp_path / "models"),
           "output_path": str(tmp_path / "rollouts"), "run_name": "t", "layers": 2, "hidden_dim": 64, "dim": 2,
           "particle_type_embedding_size": 16, "input_sequence_length": s, "connection_radius": 15.0,
           "batch_size": 2, "noise_std": 0.05, "ntraining_steps": 6, "nsave_steps": 3,
           "loss_weight_position": 1.0, "loss_weight_strain": 1.0, "lr_init": 1e-3, "lr_decay": 0.1,
           "lr_decay_steps": 100, "model_file": None, "train_state_file": None,
           "inference_mode": "autoregressive"}
    dev = torch.device("cuda")
    torch.manual_seed(0)
    sim = T._get_simulator(md, cfg["noise_std"], cfg["noise_std"], dev, cfg)
    before = {k: v.detach().cpu().clone() for k, v in sim.state_dict().items()}
    step, lowest = T.train(sim, md, dev, cfg, log_every=3)
    assert step == 6 and np.isfinite(lowest)
    assert any(not torch.equal(before[k], v.detach().cpu()) for k, v in sim.state_dict().items())
    saved = sorted((tmp_path / "models" / "t").glob("model-best-*.pt"))
    assert saved and saved[0].name == "model-best-000003.pt"
    st = torch.load(str(saved[-1]).replace("model-best", "train_state-best"))
    assert set(st) == {"optimizer_state", "global_train_state"}
    assert st["global_train_state"]["step"] == int(saved[-1].stem[-6:])

    # resume continues from the restored step
    cfg2 = dict(cfg, model_file=saved[-1].name, train_state_file=saved[-1].name.replace("model-best", "train_state-best"),
                ntraining_steps=int(saved[-1].stem[-6:]) + 2, nsave_steps=1000)
    sim2 = T._get_simulator(md, cfg["noise_std"], cfg["noise_std"], dev, cfg2)
    step2, _ = T.train(sim2, md, dev, cfg2)
    assert step2 == cfg2["ntraining_steps"]

    # rollout mode: one pickle per test case with the reference's keys
    cfg3 = dict(cfg, mode="rollout", model_file=saved[-1].name)
    sim3 = T._get_simulator(md, cfg["noise_std"], cfg["noise_std"], dev, cfg3)
    loss = T.predict(sim3, md, dev, cfg3)
    assert np.isfinite(loss)
    with open(tmp_path / "rollouts" / "t" / "case_2.pkl", "rb") as f:
        out = pickle.load(f)
    for k in ("initial_positions", "predicted_rollout", "ground_truth_rollout", "particle_types",
              "rmse_position", "rmse_strain", "metadata", "case_name"):
        assert k in out, k
    assert out["predicted_rollout"].shape == (6, 72, 2)


def test_cli_main(tmp_path):
    import yaml
    from sgnn_b200.single_scale import train as T
    s = 6
    _dataset(tmp_path / "data", s)
    cfg = {"mode": "train", "data_path": str(tmp_path / "data"), "model_path": str(tmp_path / "models"),
           "output_path": str(tmp_path / "rollouts"), "run_name": "cli", "layers": 1, "hidden_dim": 64, "dim": 2,
           "particle_type_embedding_size": 16, "input_sequence_length": s, "connection_radius": 15.0,
           "batch_size": 2, "noise_std": 0.05, "ntraining_steps": 2, "nsave_steps": 2, "lr_init": 1e-3,
           "lr_decay": 0.1, "lr_decay_steps": 100, "model_file": None, "log": False}
    p = tmp_path / "config.yaml"
    p.write_text(yaml.safe_dump(cfg))
    T.main(["--config", str(p), "--mode", "train"])
    assert (tmp_path / "models" / "cli" / "model-best-000002.pt").exists()
    T.main(["--config", str(p), "--mode", "valid", "--model_file", "model-best-000002.pt"])
