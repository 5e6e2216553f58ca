"""Device-resident multi-scale rollout with the reference's interface and output dictionary
(sgnn/multi_scale/multi_scale_evaluate.py:139-250).

The reference syncs four times per step (.item() x2, .cpu().numpy() x2, :186-196) and re-allocates
the window with torch.cat.  Here the rollout reuses the single-scale RolloutPlan: one
particle-major trajectory buffer, the decoder kernel writes frame k+S in place, all steps captured
in one CUDA graph; per-step RMSE is computed on the device after the loop and read back once.
"""
import time
from typing import Any, Dict

import numpy as np
import torch

from .. import _lib as L
from ..single_scale.evaluate import get_plan
from .multi_scale_simulator import MultiScaleSimulator


@torch.no_grad()
def evaluate_multi_scale_rollout(simulator: MultiScaleSimulator, positions: torch.Tensor,
                                 particle_type: torch.Tensor, n_particles_per_example: torch.Tensor,
                                 strains: torch.Tensor, nsteps: int, dim: int, device,
                                 input_sequence_length: int, inference_mode: str = "autoregressive",
                                 use_cuda_graph: bool = True) -> Dict[str, Any]:
    """positions [N, T, dim], strains [T, N]; rolls `nsteps` steps from the first S frames."""
    if inference_mode not in ("autoregressive", "one_step"):
        raise ValueError(f"Unknown inference_mode: {inference_mode}. Must be 'autoregressive' or 'one_step'")
    L.require_cuda(positions, "positions")
    s = input_sequence_length
    n = positions.shape[0]
    dev = positions.device
    positions = positions.float()
    t = s + nsteps
    if positions.shape[1] < t:
        raise ValueError(f"positions hold {positions.shape[1]} frames, need {t}")
    plan = get_plan(simulator, n, t, s, dim, inference_mode, dev, use_cuda_graph)
    plan.types.copy_(particle_type.to(dev))
    if inference_mode == "autoregressive":
        plan.traj[:, :s].copy_(positions[:, :s])
    else:
        plan.traj.copy_(positions[:, :t])
    plan.run()
    pred = plan.pred[:, s:t].permute(1, 0, 2)                       # [nsteps, N, dim]
    gt = positions[:, s:t].permute(1, 0, 2)
    gt_strain = strains[s:t].to(dev).float()
    rmse_pos = torch.sqrt(torch.mean(torch.sum((pred - gt) ** 2, dim=-1), dim=1))       # :186-190
    rmse_strain = torch.sqrt(torch.mean((plan.strain - gt_strain) ** 2, dim=1))
    return {
        "initial_positions": positions[:, :s].permute(1, 0, 2).cpu().numpy(),
        "initial_strains": strains[:s].cpu().numpy(),
        "predicted_rollout": pred.contiguous().cpu().numpy(),
        "ground_truth_rollout": gt.contiguous().cpu().numpy(),
        "ground_truth_strain": gt_strain.cpu().numpy(),
        "predicted_strain": plan.strain.cpu().numpy(),
        "particle_types": particle_type.cpu().numpy(),
        "rmse_position": rmse_pos.cpu().numpy().astype(np.float64),
        "rmse_strain": rmse_strain.cpu().numpy().astype(np.float64),
        "run_time": 0.0,                                            # set by the caller (:243)
        "inference_mode": inference_mode,
    }


def validate_multi_scale_simulator(simulator: MultiScaleSimulator, data_path: str, metadata: Dict[str, Any], device,
                                   input_sequence_length: int = 3, inference_mode: str = "autoregressive",
                                   num_scales: int = 3, window_size: int = 3, radius_multiplier: float = 2.0,
                                   verbose: bool = True) -> Dict[str, float]:
    """Rollout every trajectory of `data_path` with its own static graph and average the final-step
    losses (reference multi_scale_evaluate.py:18-136; same metric keys)."""
    from .static_graph_data_loader import MultiScaleTaylorImpactTrajectoriesDataset
    dataset = MultiScaleTaylorImpactTrajectoriesDataset(data_path, num_scales=num_scales, window_size=window_size,
                                                        radius_multiplier=radius_multiplier, device=device)
    was_training = simulator.training
    simulator.eval()
    simulator.to(device)
    total, position, strain, one_step, times = [], [], [], [], []
    nsteps = metadata["sequence_length"] - input_sequence_length
    for i, traj in enumerate(dataset):
        simulator.set_static_graph(traj["graph"])
        torch.cuda.synchronize()
        t0 = time.time()
        out = evaluate_multi_scale_rollout(simulator, traj["positions"], traj["particle_type"],
                                           traj["n_particles_per_example"], traj["strains"], nsteps,
                                           metadata["dim"], device, input_sequence_length, inference_mode)
        times.append(time.time() - t0)
        position.append(float(out["rmse_position"][-1]))
        strain.append(float(out["rmse_strain"][-1]))
        total.append(position[-1] + strain[-1])
        one_step.append(float(out["rmse_position"][0] + out["rmse_strain"][0]))
        if verbose:
            print(f"     Example {i + 1} - Loss total: {total[-1]:.6f}, Position: {position[-1]:.6f}, "
                  f"Strain: {strain[-1]:.6f}, Time: {times[-1]:.2f}s")
    if was_training:
        simulator.train()
    return {"val/loss_total": float(np.mean(total)), "val/loss_position": float(np.mean(position)),
            "val/loss_strain": float(np.mean(strain)), "val/loss_oneStep": float(np.mean(one_step)),
            "val/mean_time": float(np.mean(times)), "val/std_time": float(np.std(times))}
