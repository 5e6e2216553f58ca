"""Device-resident multi-scale rollout with the reference's interface and output dictionary
(sgnn/multi_scale/multi_scale_evaluate.py:139-250).

The reference syncs four times per step (.item() x2, .cpu().numpy() x2, :186-196) and re-allocates
the window with torch.cat.  Here the rollout reuses the single-scale RolloutPlan: one
particle-major trajectory buffer, the decoder kernel writes frame k+S in place, all steps captured
in one CUDA graph; per-step RMSE is computed on the device after the loop and read back once.
"""
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
