"""Rollout inference of the multi-scale simulator over a trajectory file (reference
sgnn/multi_scale/multi_scale_inference.py:19-129): load the checkpoint, roll every trajectory of the
valid (mode 'valid') or test (mode 'rollout') split out with ITS static graph, print the per-case
losses and, in rollout mode, write one pickle per case under output_path/run_name."""
import pickle
import time
from pathlib import Path

import torch

from . import multi_scale_evaluate
from .static_graph_data_loader import get_multi_scale_data_loader_by_trajectories


def run_inference(simulator, metadata, config, device):
    """Config keys read (the reference's, :22-26): model_path, run_name, model_file, output_path, data_path,
    input_sequence_length, inference_mode, dim, num_scales, window_size, radius_multiplier, mode.
    Returns the mean over cases of final rmse_position + rmse_strain (the reference only prints it)."""
    model_path = Path(config["model_path"]) / config["run_name"] / config["model_file"]
    simulator.load(str(model_path))           # a missing / unreadable checkpoint raises (the reference exits, :33-42)
    simulator.to(device)
    simulator.eval()
    split = "test" if config["mode"] == "rollout" else "valid"
    trajectories = get_multi_scale_data_loader_by_trajectories(
        path=str(Path(config["data_path"]) / f"{split}.npz"), num_scales=config["num_scales"],
        window_size=config["window_size"], radius_multiplier=config["radius_multiplier"], device=device)
    nsteps = metadata["sequence_length"] - config["input_sequence_length"]
    names = metadata.get("file_test" if split == "test" else "file_valid") or []
    eval_loss, total_time = [], 0.0
    with torch.no_grad():
        for i, traj in enumerate(trajectories):
            simulator.set_static_graph(traj["graph"])
            torch.cuda.synchronize()
            t0 = time.time()
            out = multi_scale_evaluate.evaluate_multi_scale_rollout(
                simulator=simulator, positions=traj["positions"], particle_type=traj["particle_type"],
                n_particles_per_example=traj["n_particles_per_example"], strains=traj["strains"], nsteps=nsteps,
                dim=config["dim"], device=device, input_sequence_length=config["input_sequence_length"],
                inference_mode=config["inference_mode"])
            out["run_time"] = time.time() - t0
            total_time += out["run_time"]
            out["metadata"] = metadata
            loss_position, loss_strain = float(out["rmse_position"][-1]), float(out["rmse_strain"][-1])
            name = names[i] if i < len(names) else f"rollout_{i}.npz"
            print(f"Predicting example {i}-\n      {name}\n      loss_total: {loss_position + loss_strain}, "
                  f"loss_position: {loss_position}, loss_strain: {loss_strain}\n  Runtime: {out['run_time']:.2f}s")
            eval_loss.append(loss_position + loss_strain)
            if config["mode"] == "rollout":
                out["case_name"] = name.replace(".npz", "")
                save_dir = Path(config["output_path"]) / config["run_name"]
                save_dir.mkdir(parents=True, exist_ok=True)
                with open(save_dir / f"{out['case_name']}.pkl", "wb") as f:
                    pickle.dump(out, f)
    mean = sum(eval_loss) / max(1, len(eval_loss))
    print(f"Mean loss: {mean:.6f}\nTotal runtime: {total_time:.2f}s\n"
          f"Average runtime per rollout: {total_time / max(1, len(eval_loss)):.2f}s\n"
          f"Peak VRAM usage: {torch.cuda.max_memory_allocated(device) / 2 ** 20:.1f}MB")
    return mean
