"""Device-resident rollout driver with the reference's interface and output dictionary
(sgnn/single_scale/evaluate.py:23-174).

The reference re-allocates the [N,S,dim] window with torch.cat every step and syncs twice per
step inside the neighbour search.  Here the trajectory lives in ONE particle-major buffer
[N, S+nsteps, dim]; step k reads frames [k, k+S) in place and the decoder kernel writes frame
k+S, so there is no window shifting at all, and the whole rollout (nsteps x ~26 kernels) is
captured once into a CUDA graph and replayed with a single launch.
"""
import time

import numpy as np
import torch

from .. import _lib as L
from ..blob import params_version
from .learned_simulator import LearnedSimulator, example_ptr_of

EROSIONAL_PARTICLE_ID = -1      # evaluate.py:16
RMSE_PRINT_INTERVAL = 3         # evaluate.py:20
VERBOSE = True
MAX_CACHED_PLANS = 4            # each plan owns a trajectory buffer, a CUDA graph and the engine it captured


def rollout_rmse(pred: np.ndarray, gt: np.ndarray) -> np.ndarray:
    """Accumulated RMSE over time, shape (timesteps,) (reference rollout_rmse, evaluate.py:23-48)."""
    if pred.shape != gt.shape:
        raise ValueError(f"Shape mismatch: pred {pred.shape} vs gt {gt.shape}")
    t = gt.shape[0]
    sq = np.square(pred - gt).reshape(t, -1)
    loss = np.sqrt(np.cumsum(np.mean(sq, axis=1), axis=0) / np.arange(1, t + 1))
    if VERBOSE:
        for step in range(0, t, RMSE_PRINT_INTERVAL):
            print("Testing rmse @ step %d loss: %.2e" % (step, loss[step]))
    return loss


class RolloutPlan:
    """Static buffers + captured CUDA graph of one (N, T, S, mode) rollout shape."""

    def __init__(self, sim: LearnedSimulator, n: int, total_frames: int, s: int, dim: int,
                 mode: str, device, use_graph: bool = True, n_examples: int = 1):
        self.sim, self.n, self.t, self.s, self.dim, self.mode = sim, n, total_frames, s, dim, mode
        self.nsteps = total_frames - s
        f32 = dict(dtype=torch.float32, device=device)
        self.traj = torch.zeros(n, total_frames, dim, **f32)       # window source
        # one_step: predictions go to their own buffer (the window keeps reading ground truth)
        self.pred = self.traj if mode == "autoregressive" else torch.zeros(n, total_frames, dim, **f32)
        self.strain = torch.zeros(self.nsteps, n, **f32)
        self.types = torch.zeros(n, dtype=torch.int64, device=device)
        self.example_ptr = example_ptr_of([n], device)             # one example; B + 1 boundaries for a batch of B
        if n_examples > 1:                                          # (filled by the caller: the graph reads the buffer)
            self.example_ptr = torch.zeros(n_examples + 1, dtype=torch.int32, device=device)
        self.graph = None
        self.use_graph = use_graph
        self.version = None
        self.launches_per_replay = 0
        self.captured = None      # (engine, edge set(s)) whose device pointers are baked into `graph`

    def launch_steps(self, lo: int, hi: int):
        used = None
        for k in range(lo, hi):
            used = self.sim._step(self.traj[:, k:k + self.s], None, self.types,
                                  next_pos=self.pred[:, k + self.s], strain=self.strain[k],
                                  example_ptr=self.example_ptr)
        return used

    def run(self):
        """All steps on the current stream (graph replay when captured)."""
        if not self.use_graph:
            self.launch_steps(0, self.nsteps)
            return
        ver = params_version(self.sim.parameters())
        if self.graph is None or self.version != ver:
            self.sim._prepare(self.n, self.traj.device)            # (re)pack weights outside capture
            saved = self.traj[:, self.s].clone()
            self.launch_steps(0, 1)                                # warm-up: func attributes, lazy init
            torch.cuda.synchronize()
            if self.mode == "autoregressive":
                self.traj[:, self.s].copy_(saved)
            g = torch.cuda.CUDAGraph()
            before = L.launch_count
            with torch.cuda.graph(g):
                used = self.launch_steps(0, self.nsteps)
            self.launches_per_replay = L.launch_count - before
            # The graph replays raw pointers into the engine's weight blobs / latent buffers and the edge-set
            # buffers.  The simulator's engine cache is bounded and evicts; the plan therefore OWNS what it
            # captured, so a replay can never touch freed or re-used allocator memory.
            self.graph, self.version, self.captured = g, ver, used
        self.graph.replay()
        L.launch_count += self.launches_per_replay


def get_plan(sim, n, t, s, dim, mode, device, use_graph=True, n_examples: int = 1) -> RolloutPlan:
    cache = sim.__dict__.setdefault("_rollout_plans", {})
    key = (n, t, s, dim, mode, str(device), use_graph, getattr(sim, "_graph_token", 0), n_examples)
    plan = cache.pop(key, None)
    if plan is None:
        plan = RolloutPlan(sim, n, t, s, dim, mode, device, use_graph, n_examples)
        while len(cache) >= MAX_CACHED_PLANS:          # least recently used first (dicts keep insertion order)
            cache.pop(next(iter(cache)))
    cache[key] = plan
    return plan


@torch.no_grad()
def rollout(simulator: LearnedSimulator, position: torch.Tensor, particle_types: torch.Tensor,
            n_particles_per_example: torch.Tensor, strains: torch.Tensor, nsteps: int,
            particle_dim: int, device, input_sequence_length: int = 3,
            inference_mode: str = "autoregressive", use_cuda_graph: bool = True) -> dict:
    """Rolls a trajectory out by applying the model recursively (reference rollout, evaluate.py:52).
    position [N, T, dim], strains [T, N]; as in the reference, nsteps is taken from the data."""
    if position.dim() != 3:
        raise ValueError(f"Position tensor must be 3D, got {position.dim()}D")
    if strains.dim() != 2:
        raise ValueError(f"Strains tensor must be 2D, got {strains.dim()}D")
    if position.shape[0] != strains.shape[1]:
        raise ValueError(f"Number of particles mismatch: position {position.shape[0]} vs strains {strains.shape[1]}")
    if position.shape[1] < input_sequence_length:
        raise ValueError(f"Position sequence length {position.shape[1]} must be >= input_sequence_length {input_sequence_length}")
    if inference_mode not in ("autoregressive", "one_step"):
        raise ValueError(f"Unknown inference_mode: {inference_mode}. Must be 'autoregressive' or 'one_step'")
    L.require_cuda(position, "position")
    s = input_sequence_length
    n, t, dim = position.shape
    nsteps = t - s                                     # evaluate.py:106
    dev = position.device
    position = position.float()
    particle_types = particle_types.to(dev)
    erosional = particle_types == EROSIONAL_PARTICLE_ID
    has_erosional = bool(erosional.any().item())        # one sync, before the timed loop
    # a [B] tensor of particle counts (B > 1) rolls B independent trajectories out as one system: examples do not
    # interact (the neighbour search is per example, learned_simulator.py:104-117 of the reference)
    n_examples = n_particles_per_example.numel() if isinstance(n_particles_per_example, torch.Tensor) else 1
    plan = get_plan(simulator, n, t, s, dim, inference_mode, dev, use_cuda_graph and not has_erosional,
                    max(n_examples, 1))
    if n_examples > 1:
        plan.example_ptr.copy_(example_ptr_of(n_particles_per_example, dev))
    plan.types.copy_(particle_types)
    gt_strains = strains[s:].to(dev).float()

    if inference_mode == "autoregressive":
        plan.traj[:, :s].copy_(position[:, :s])
    else:
        plan.traj.copy_(position)
    torch.cuda.synchronize()
    start = time.time()
    if not has_erosional:
        plan.run()
    else:
        # prescribed (broken) particles follow the ground truth (evaluate.py:113-131)
        for k in range(nsteps):
            plan.launch_steps(k, k + 1)
            frame = plan.pred[:, k + s]
            frame.copy_(torch.where(erosional[:, None], position[:, k + s], frame))
            plan.strain[k].copy_(torch.where(erosional, gt_strains[k], plan.strain[k]))
    torch.cuda.synchronize()
    run_time = time.time() - start

    # Metrics on the device, then every output in ONE batch of asynchronous copies into pinned host
    # memory (torch's caching host allocator: fresh blocks per call, so results of earlier calls are
    # never overwritten) and a single synchronisation.  The reference does the RMSE in numpy after
    # four separate pageable copies (evaluate.py:147-160).
    pred_positions = plan.pred[:, s:].permute(1, 0, 2)
    gt_positions = position[:, s:].permute(1, 0, 2)
    dev_out = {
        "initial_positions": position[:, :s].permute(1, 0, 2),
        "initial_strains": strains[:s],
        "predicted_rollout": pred_positions,
        "ground_truth_rollout": gt_positions,
        "ground_truth_strain": gt_strains,
        "predicted_strain": plan.strain,
        "particle_types": particle_types,
        "rmse_position": _rollout_rmse_device(pred_positions, gt_positions),
        "rmse_strain": _rollout_rmse_device(plan.strain, gt_strains),
    }
    host = {}
    for k, v in dev_out.items():
        h = torch.empty(v.shape, dtype=v.dtype, pin_memory=True)
        h.copy_(v, non_blocking=True)
        host[k] = h
    torch.cuda.synchronize()
    out = {k: h.numpy() for k, h in host.items()}
    for name in ("rmse_position", "rmse_strain"):
        _print_rmse(out[name])
    out["run_time"] = run_time
    out["inference_mode"] = inference_mode
    return out


def _rollout_rmse_device(pred: torch.Tensor, gt: torch.Tensor) -> torch.Tensor:
    """rollout_rmse (evaluate.py:23-48) on the device: sqrt(cumsum_t(mean_n (pred - gt)^2) / (t + 1))."""
    t = gt.shape[0]
    sq = (pred.float() - gt.float()).square().reshape(t, -1).mean(dim=1)
    return torch.sqrt(torch.cumsum(sq, dim=0) / torch.arange(1, t + 1, device=sq.device, dtype=sq.dtype))


def _print_rmse(loss: np.ndarray):
    if VERBOSE:
        for step in range(0, loss.shape[0], RMSE_PRINT_INTERVAL):
            print("Testing rmse @ step %d loss: %.2e" % (step, loss[step]))
