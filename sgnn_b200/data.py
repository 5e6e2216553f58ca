"""Data path of the train step and of the rollout evaluation.

File format = the reference's (datasets/taylor_impact_2d/build_dataset.py:313-346, loader
taylor_impact_data_loader.py:69-93): an NPZ with one pickled dict `trajectories`,
{name: (positions [T,N,dim] float, particle_types [N] int, stresses [T,N] float)}, and a
`metadata.json` beside it.  Batches have the layout of the reference's `collate_fn`
(taylor_impact_data_loader.py:243-284): {'input': {positions [sum N,S,dim], particle_type [sum N],
n_particles_per_example [B]}, 'output': {next_position [sum N,dim], next_strain [sum N]},
'meta': {trajectory_idx [B], time_idx [B]}}.

B200 design: the reference slices numpy arrays per sample, vstacks them on the host and copies five
tensors to the device every step.  A dataset of this kind (hundreds of trajectories x hundreds of
frames x thousands of particles x dim floats: a few GB) fits in a corner of 180 GB of HBM, so the
trajectories are uploaded ONCE and a batch is a handful of device-side slices: no host work and no
H2D copy inside the training loop.  Data-parallel ranks draw disjoint slices of the same seeded
permutation.
"""
import json
from pathlib import Path
from typing import Dict, Iterator, List, Optional, Tuple

import numpy as np
import torch

Trajectory = Tuple[np.ndarray, np.ndarray, np.ndarray]


def load_trajectories(data_path: str) -> List[Trajectory]:
    """[(positions [T,N,dim], particle_types [N], stresses [T,N])] in file order; entries that are
    not 3-tuples are skipped with a warning (taylor_impact_data_loader.py:83-93)."""
    p = Path(data_path)
    if not p.exists():
        raise FileNotFoundError(f"Data file not found: {p}")
    with np.load(p, allow_pickle=True) as f:
        trajectories = f["trajectories"].item()
    out = []
    for name, tr in trajectories.items():
        if isinstance(tr, tuple) and len(tr) == 3:
            out.append(tr)
        else:
            print(f"Warning: Skipping trajectory {name} with unexpected format")
    return out


def save_trajectories(data_path: str, trajectories: Dict[str, Trajectory], metadata: Optional[dict] = None):
    """Write the reference's NPZ layout (+ metadata.json beside it when given)."""
    p = Path(data_path)
    p.parent.mkdir(parents=True, exist_ok=True)
    np.savez_compressed(p, trajectories=np.array(trajectories, dtype=object))
    if metadata is not None:
        with open(p.parent / "metadata.json", "w") as f:
            json.dump(metadata, f)


def load_metadata(data_dir: str) -> dict:
    with open(Path(data_dir) / "metadata.json") as f:
        return json.load(f)


def stress_stats(data_path: str) -> Optional[Dict[str, float]]:
    """{'mean','std'} from metadata.json next to the NPZ, or None (taylor_impact_data_loader.py:20-52)."""
    m = Path(data_path).parent / "metadata.json"
    if not m.exists():
        return None
    with open(m) as f:
        md = json.load(f)
    if md.get("stress_mean") is None or md.get("stress_std") is None:
        return None
    return {"mean": md["stress_mean"], "std": md["stress_std"]}


def compute_metadata(trajectories: List[Trajectory], extra: Optional[dict] = None) -> dict:
    """vel/acc mean and std over all frames (the keys train._get_simulator reads, train.py:446-459)."""
    vs, as_ = [], []
    for pos, _, _ in trajectories:
        v = np.diff(np.asarray(pos, dtype=np.float64), axis=0)
        vs.append(v.reshape(-1, v.shape[-1]))
        a = np.diff(v, axis=0)
        as_.append(a.reshape(-1, a.shape[-1]))
    v, a = np.concatenate(vs), np.concatenate(as_)
    md = {"dim": int(v.shape[-1]), "sequence_length": int(trajectories[0][0].shape[0]),
          "num_particle_types": 1,
          "vel_mean": v.mean(0).tolist(), "vel_std": v.std(0).tolist(),
          "acc_mean": a.mean(0).tolist(), "acc_std": a.std(0).tolist()}
    md.update(extra or {})
    return md


class DeviceSamples:
    """All one-step training samples of a trajectory file, resident on the device.

    Sample idx -> (trajectory, time) exactly as TaylorImpactSamplesDataset.__getitem__
    (taylor_impact_data_loader.py:143-181): trajectories contribute T - S samples each, time index
    S + (idx - start); the input window is frames [t-S, t) transposed to [N,S,dim], the targets are
    frame t; every particle gets particle_types[0] (the reference's np.full, :161)."""

    def __init__(self, data_path_or_list, input_length_sequence: int = 6, device="cuda"):
        trs = load_trajectories(data_path_or_list) if isinstance(data_path_or_list, (str, Path)) \
            else list(data_path_or_list)
        if len(trs) == 0:
            raise ValueError(f"No trajectories found in {data_path_or_list}")
        self.device = torch.device(device)
        self.s = int(input_length_sequence)
        self.pos = [torch.as_tensor(np.ascontiguousarray(p, dtype=np.float32)).to(self.device) for p, _, _ in trs]
        self.stress = [torch.as_tensor(np.ascontiguousarray(q, dtype=np.float32)).to(self.device) for _, _, q in trs]
        self.ptype = [int(np.asarray(t).reshape(-1)[0]) for _, t, _ in trs]
        self.dim = int(self.pos[0].shape[-1])
        lengths = [int(p.shape[0]) - self.s for p in self.pos]
        if min(lengths) < 1:
            raise ValueError("a trajectory is shorter than input_length_sequence + 1 frames")
        self.starts = np.cumsum([0] + lengths[:-1])
        self.length = int(sum(lengths))

    def __len__(self):
        return self.length

    def locate(self, idx: int) -> Tuple[int, int]:
        tr = int(np.searchsorted(self.starts, idx, side="right") - 1)
        return tr, self.s + (int(idx) - int(self.starts[tr]))

    def batch(self, indices) -> Dict[str, Dict[str, torch.Tensor]]:
        """The reference's collated batch, assembled from device slices."""
        where = [self.locate(int(i)) for i in indices]
        dev = self.device
        windows = [self.pos[tr][t - self.s:t].permute(1, 0, 2) for tr, t in where]
        counts = [w.shape[0] for w in windows]
        types = torch.cat([torch.full((n,), self.ptype[tr], dtype=torch.int64, device=dev)
                           for (tr, _), n in zip(where, counts)])
        return {
            "input": {"positions": torch.cat(windows).contiguous(), "particle_type": types,
                      "n_particles_per_example": torch.tensor(counts, dtype=torch.int64, device=dev)},
            "output": {"next_position": torch.cat([self.pos[tr][t] for tr, t in where]).contiguous(),
                       "next_strain": torch.cat([self.stress[tr][t] for tr, t in where]).contiguous()},
            "meta": {"trajectory_idx": torch.tensor([tr for tr, _ in where]),
                     "time_idx": torch.tensor([t for _, t in where])},
        }


class DeviceSampleLoader:
    """Iterates batches of `batch_size` samples PER RANK.  One epoch = one seeded permutation of all
    samples shared by every rank (same seed), of which rank r takes the r-th slice of each global
    batch of world*batch_size samples; the ragged tail of an epoch is dropped on multi-rank runs so
    that every rank takes the same number of steps (the allreduce would hang otherwise)."""

    def __init__(self, samples: DeviceSamples, batch_size: int = 2, shuffle: bool = True, seed: int = 0,
                 rank: int = 0, world: int = 1):
        self.samples, self.batch_size, self.shuffle = samples, int(batch_size), shuffle
        self.rank, self.world, self.seed, self.epoch = rank, world, seed, 0

    def __len__(self):
        g = self.batch_size * self.world
        n = len(self.samples)
        return n // g if self.world > 1 else (n + g - 1) // g

    def __iter__(self) -> Iterator[Dict]:
        n = len(self.samples)
        if self.shuffle:
            gen = torch.Generator().manual_seed(self.seed + self.epoch)
            order = torch.randperm(n, generator=gen).tolist()
        else:
            order = list(range(n))
        self.epoch += 1
        g = self.batch_size * self.world
        for b in range(len(self)):
            chunk = order[b * g:(b + 1) * g]
            mine = chunk[self.rank * self.batch_size:(self.rank + 1) * self.batch_size]
            if mine:
                yield self.samples.batch(mine)


def get_data_loader_by_samples(path: str, input_length_sequence: int = 6, batch_size: int = 2,
                               shuffle: bool = True, device="cuda", seed: int = 0, rank: int = 0,
                               world: int = 1, **_ignored) -> DeviceSampleLoader:
    """Same name and leading arguments as taylor_impact_data_loader.py:287-320 (num_workers /
    pin_memory are accepted and ignored: there is no host stage to parallelise or pin)."""
    return DeviceSampleLoader(DeviceSamples(path, input_length_sequence, device), batch_size, shuffle,
                              seed, rank, world)


class TrajectoriesOnDevice:
    """Whole trajectories for rollout evaluation, items as TaylorImpactTrajectoriesDataset
    (taylor_impact_data_loader.py:219-240): positions [N,T,dim], particle_type [N],
    n_particles_per_example 0-d, strains [T,N], trajectory_idx."""

    def __init__(self, data_path_or_list, device="cuda"):
        self.trs = load_trajectories(data_path_or_list) if isinstance(data_path_or_list, (str, Path)) \
            else list(data_path_or_list)
        self.device = torch.device(device)

    def __len__(self):
        return len(self.trs)

    def __getitem__(self, idx: int) -> Dict:
        pos, types, stress = self.trs[idx]
        p = torch.as_tensor(np.ascontiguousarray(pos, dtype=np.float32)).to(self.device)
        n = p.shape[1]
        return {"positions": p.permute(1, 0, 2).contiguous(),
                "particle_type": torch.full((n,), int(np.asarray(types).reshape(-1)[0]), dtype=torch.int64,
                                            device=self.device),
                "n_particles_per_example": torch.tensor(n, device=self.device),
                "strains": torch.as_tensor(np.ascontiguousarray(stress, dtype=np.float32)).to(self.device),
                "trajectory_idx": idx}

    def __iter__(self):
        return (self[i] for i in range(len(self)))


def get_data_loader_by_trajectories(path: str, device="cuda", **_ignored) -> TrajectoriesOnDevice:
    return TrajectoriesOnDevice(path, device)


def synthetic_taylor_trajectory(nx: int = 120, ny: int = 40, spacing: float = 12.5, frames: int = 32,
                                seed: int = 0) -> Trajectory:
    """The SURVEY.md 8(d) synthetic Taylor bar as one trajectory in the file layout."""
    g = torch.Generator().manual_seed(seed)
    xs = (torch.arange(nx, dtype=torch.float64) + 1.0) * spacing - 2.0
    ys = (torch.arange(ny, dtype=torch.float64) - (ny - 1) / 2.0) * spacing
    base = torch.stack(torch.meshgrid(xs, ys, indexing="ij"), dim=-1).reshape(-1, 2)
    n = base.shape[0]
    pos = torch.empty(frames, n, 2, dtype=torch.float64)
    for t in range(frames):
        p = base.clone()
        p[:, 0] += -0.25 * spacing * t - 0.02 * spacing * t * torch.exp(-(base[:, 0] + 2.0) / (10 * spacing))
        if t > 0:
            p += 0.01 * spacing * torch.randn(n, 2, generator=g, dtype=torch.float64)
        pos[t] = p
    stress = torch.randn(frames, n, generator=g)
    return pos.float().numpy(), np.zeros(n, dtype=np.int64), stress.numpy()
