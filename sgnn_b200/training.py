"""Training step of the learned simulators (reference sgnn/single_scale/train.py:230-280 and
sgnn/multi_scale/multi_scale_train.py:130-190) with data parallelism over NCCL/NVLink.

The reference is single-process.  Here every rank owns a shard of the batch (whole samples: they
are disjoint graphs, learned_simulator.py:104-117); the loss is the mean over ALL particles of the
global batch (train.py:267-268), so each rank scales its local sum by 1/N_global and the fp32
gradients are SUM-allreduced as ONE flat buffer (0.64 MB at H=64: latency-bound, a single NCCL
launch); every rank then applies the identical Adam update.  world_size 1 skips the collective.
"""
from typing import Dict, Optional

import torch
import torch.distributed as dist

from . import noise_utils


def learning_rate(step: int, lr_init: float, lr_decay: float, lr_decay_steps: float) -> float:
    """lr(step) of the reference, set AFTER the optimizer step (train.py:276-278)."""
    return lr_init * (lr_decay ** (step / lr_decay_steps)) + 1e-6


def loss_terms(pred_acc, target_acc, pred_strain, next_strain, w_pos: float = 1.0, w_strain: float = 1.0):
    """Per-particle loss w_p * sum_d (a^ - a)^2 + w_s (s^ - s)^2 (train.py:257-266)."""
    return w_pos * ((pred_acc - target_acc) ** 2).sum(dim=-1) + w_strain * (pred_strain - next_strain) ** 2


class FlatGradAllReduce:
    """SUM-allreduce of all gradients as one flat fp32 buffer.  Parameters that received no
    gradient (the unused particle-type embedding when there is one type) contribute zeros so that
    every rank reduces the same layout."""

    def __init__(self, params, group=None):
        self.params = [p for p in params if p.requires_grad]
        self.group = group
        self.numel = sum(p.numel() for p in self.params)
        self.flat = None

    def __call__(self):
        if not dist.is_initialized() or dist.get_world_size(self.group) == 1:
            return
        p0 = self.params[0]
        if self.flat is None or self.flat.device != p0.device:
            self.flat = torch.zeros(self.numel, dtype=torch.float32, device=p0.device)
        off = 0
        for p in self.params:
            n = p.numel()
            if p.grad is None:
                self.flat[off:off + n].zero_()
            else:
                self.flat[off:off + n].copy_(p.grad.reshape(-1))
            off += n
        dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)
        off = 0
        for p in self.params:          # hand the reduced buffer out as views: no copy back
            n = p.numel()
            p.grad = self.flat[off:off + n].view_as(p)
            off += n


def global_particle_count(n_local: int, device, group=None) -> int:
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return n_local
    t = torch.tensor([n_local], dtype=torch.int64, device=device)
    dist.all_reduce(t, group=group)
    return int(t.item())


def make_optimizer(params, lr: float, device=None) -> torch.optim.Adam:
    """Adam as the reference builds it (train.py:199) in the form a captured step needs: fused, capturable,
    learning rate held in a device tensor (set_lr updates it in place)."""
    params = list(params)
    dev = params[0].device if device is None else device
    return torch.optim.Adam(params, lr=torch.tensor(float(lr), dtype=torch.float32, device=dev), fused=True,
                            capturable=True)


def set_lr(optimizer, lr: float):
    for g in optimizer.param_groups:
        if isinstance(g["lr"], torch.Tensor):
            g["lr"].fill_(float(lr))
        else:
            g["lr"] = float(lr)


class matmul_precision:
    """Scope for the dense GEMMs of the training path (cuBLAS through torch).  "fp32" (default) is what
    the reference runs and what the gradient parity tests pin; "tf32" lets cuBLAS use the tensor cores
    (TF32 operands, fp32 accumulate -- the precision the inference kernels use)."""

    def __init__(self, mode: str):
        if mode not in ("fp32", "tf32"):
            raise ValueError("matmul precision must be 'fp32' or 'tf32'")
        self.mode = mode

    def __enter__(self):
        self.saved = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = self.mode == "tf32"

    def __exit__(self, *a):
        torch.backends.cuda.matmul.allow_tf32 = self.saved


def train_step(simulator, optimizer, batch: Dict[str, torch.Tensor], noise_std: float,
               w_pos: float = 1.0, w_strain: float = 1.0, reducer: Optional[FlatGradAllReduce] = None,
               n_global: Optional[int] = None, noise: Optional[torch.Tensor] = None,
               generator: Optional[torch.Generator] = None, matmul: str = "fp32", engine=None) -> torch.Tensor:
    """One optimisation step on this rank's shard.  batch: positions [N,S,dim], particle_type [N],
    n_particles_per_example [B], next_position [N,dim], next_strain [N] (the tensors the
    reference's collate_fn emits, taylor_impact_data_loader.py:243-284), already on the device.
    Returns this rank's contribution to the global mean loss (sum over ranks = the loss)."""
    if engine is not None:          # hand-written forward / backward / allreduce / Adam under one CUDA graph
        return engine.step(batch, noise_std, noise=noise, generator=generator, n_global=n_global)
    pos = batch["positions"]
    if noise is None:
        noise = noise_utils.get_random_walk_noise_for_position_sequence(pos, noise_std, generator)
    optimizer.zero_grad(set_to_none=True)
    with matmul_precision(matmul):
        pred_acc, target_acc, pred_strain = simulator.predict_accelerations(
            next_positions=batch["next_position"], position_sequence_noise=noise, position_sequence=pos,
            nparticles_per_example=batch["n_particles_per_example"], particle_types=batch["particle_type"])
        per_particle = loss_terms(pred_acc, target_acc, pred_strain, batch["next_strain"], w_pos, w_strain)
        n_glob = pos.shape[0] if n_global is None else n_global
        loss = per_particle.sum() / n_glob
        loss.backward()
    if reducer is not None:
        reducer()
    optimizer.step()
    return loss.detach()
