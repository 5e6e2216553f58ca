"""Synthetic Taylor-Impact-2D-shaped particle data (the real dataset is not available offline).

Layouts follow what the reference's loaders emit
(datasets/taylor_impact_2d/taylor_impact_data_loader.py:143-181,243-284):
  positions       [N, T, dim] float32   (particle-major, as `rollout` receives it, evaluate.py:102)
  particle_types  [N] int64             (all 0: one particle type, build_dataset.py:240)
  strains         [T, N] float32
  metadata        dict with dim, sequence_length, num_particle_types, vel/acc mean and std

Geometry and motion are the recipe in SURVEY.md section 8(d): an nx x ny bar lattice at spacing
`s`, leading edge one spacing from the rigid wall at x = -2, frame 0 exactly on the lattice;
rigid translation + smooth compressive field + small iid jitter from frame 1 on.
"""
from typing import Dict, Tuple

import numpy as np
import torch


def taylor_bar_lattice(nx: int = 120, ny: int = 40, spacing: float = 0.5) -> torch.Tensor:
    """[nx*ny, 2] float32 lattice; x-major ordering (particle id = ix * ny + iy)."""
    xs = (-2.0 + spacing * (1 + np.arange(nx, dtype=np.float64))).astype(np.float32)
    ys = (spacing * (np.arange(ny, dtype=np.float64) - (ny - 1) / 2.0)).astype(np.float32)
    xx, yy = np.meshgrid(xs, ys, indexing="ij")
    return torch.from_numpy(np.stack([xx.ravel(), yy.ravel()], axis=1).copy())


def taylor_trajectory(nx: int = 120, ny: int = 40, spacing: float = 0.5, nframes: int = 22,
                      seed: int = 0) -> Dict[str, torch.Tensor]:
    """One synthetic trajectory. Returns positions [N,T,2], particle_types [N], strains [T,N]."""
    g = torch.Generator().manual_seed(seed)
    p0 = taylor_bar_lattice(nx, ny, spacing).to(torch.float64)
    n = p0.shape[0]
    frames = []
    for t in range(nframes):
        p = p0.clone()
        if t > 0:
            p[:, 0] += -0.25 * spacing * t
            p[:, 0] += -0.02 * spacing * t * torch.exp(-(p0[:, 0] + 2.0) / (10.0 * spacing))
            p += 0.01 * spacing * torch.randn(n, 2, generator=g, dtype=torch.float64)
        frames.append(p)
    positions = torch.stack(frames, dim=1).to(torch.float32).contiguous()
    strains = torch.randn(nframes, n, generator=g, dtype=torch.float32)
    return {
        "positions": positions,
        "particle_types": torch.zeros(n, dtype=torch.int64),
        "strains": strains,
    }


def metadata_from(positions: torch.Tensor) -> Dict:
    """vel/acc statistics the way build_dataset.py derives them (per-dimension mean/std)."""
    vel = positions[:, 1:] - positions[:, :-1]
    acc = vel[:, 1:] - vel[:, :-1]
    dim = positions.shape[-1]
    return {
        "dim": dim,
        "sequence_length": positions.shape[1],
        "num_particle_types": 1,
        "vel_mean": vel.reshape(-1, dim).mean(0).tolist(),
        "vel_std": vel.reshape(-1, dim).std(0).tolist(),
        "acc_mean": acc.reshape(-1, dim).mean(0).tolist(),
        "acc_std": acc.reshape(-1, dim).std(0).tolist(),
    }


def normalization_stats(metadata: Dict, noise_std: float, device="cpu") -> Dict:
    """Same combination as the reference's `_get_simulator` (single_scale/train.py:446-457)."""
    f = lambda v: torch.tensor(v, dtype=torch.float32)  # noqa: E731
    return {
        "acceleration": {
            "mean": f(metadata["acc_mean"]).to(device),
            "std": torch.sqrt(f(metadata["acc_std"]) ** 2 + noise_std ** 2).to(device),
        },
        "velocity": {
            "mean": f(metadata["vel_mean"]).to(device),
            "std": torch.sqrt(f(metadata["vel_std"]) ** 2 + noise_std ** 2).to(device),
        },
    }


def training_batch(traj: Dict[str, torch.Tensor], input_sequence_length: int, batch: int,
                   seed: int = 0) -> Dict[str, torch.Tensor]:
    """`batch` samples cut from one trajectory, collated the way the reference does
    (taylor_impact_data_loader.py:243-284): samples are concatenated along the particle axis."""
    pos, strains = traj["positions"], traj["strains"]
    n, t, _ = pos.shape
    rng = np.random.default_rng(seed)
    starts = rng.integers(0, t - input_sequence_length, size=batch)
    S = input_sequence_length
    return {
        "positions": torch.cat([pos[:, s:s + S] for s in starts], dim=0).contiguous(),
        "particle_type": traj["particle_types"].repeat(batch),
        "n_particles_per_example": torch.full((batch,), n, dtype=torch.int64),
        "next_position": torch.cat([pos[:, s + S] for s in starts], dim=0).contiguous(),
        "next_strain": torch.cat([strains[s + S] for s in starts], dim=0).contiguous(),
    }


def spacing_for(config: str) -> Tuple[float, float]:
    """(spacing, radius) of the three synthetic connectivity regimes of SURVEY.md section 8(d)."""
    return {"A": (12.5, 15.0), "B": (5.0, 15.0), "C": (0.5, 15.0), "yaml": (0.5, 0.6)}[config]
