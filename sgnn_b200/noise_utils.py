"""Random-walk position noise for training (reference sgnn/noise_utils.py:4-39), generated on the
device of `position_sequence` (the reference draws it on the CPU and copies it over every step,
train.py:241-243)."""
import torch


def get_random_walk_noise_for_position_sequence(position_sequence: torch.Tensor, noise_std_last_step,
                                                generator: torch.Generator = None) -> torch.Tensor:
    n, s, dim = position_sequence.shape
    num_velocities = s - 1
    vel_noise = torch.randn(n, num_velocities, dim, device=position_sequence.device, generator=generator,
                            dtype=position_sequence.dtype) * (noise_std_last_step / num_velocities ** 0.5)
    vel_noise = torch.cumsum(vel_noise, dim=1)                         # random walk in velocity
    return torch.cat([torch.zeros_like(vel_noise[:, 0:1]), torch.cumsum(vel_noise, dim=1)], dim=1)
