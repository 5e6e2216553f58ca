"""Data path of the multi-scale simulator: trajectories with one STATIC multi-scale graph each
(reference sgnn/multi_scale/static_graph_data_loader.py: build_static_multi_scale_graph :27,
MultiScaleTaylorImpactSamplesDataset :63, MultiScaleTaylorImpactTrajectoriesDataset :141,
multi_scale_collate_fn :212, get_multi_scale_data_loader_by_samples :235,
get_multi_scale_data_loader_by_trajectories :281).

Same function names, arguments and item / batch dictionaries as the reference; the differences are
where the work happens:

  * the graph of a trajectory is built from its frame-0 positions on the GPU (neighbour-search
    kernel behind MultiScaleGraph) and stays on the device, as do the trajectories
    (sgnn_b200/data.py): a batch is a handful of device slices plus a reference to a cached graph;
  * batches keep the reference's collate quirk (:226-229): the batch carries the graph of its
    FIRST sample, whatever trajectories the other samples come from;
  * `rank` / `world` shard the seeded sample permutation by whole samples ("shard by trajectory
    sample"); num_workers / pin_memory are accepted and ignored.
"""
from pathlib import Path
from typing import Any, Dict, Iterator, List

import numpy as np
import torch

from .. import data as D
from .multi_scale_graph import MultiScaleConfig, MultiScaleGraph


def build_static_multi_scale_graph(initial_positions: torch.Tensor, num_scales: int = 3, window_size: int = 3,
                                   radius_multiplier: float = 2.0) -> Dict[str, Any]:
    """{'graph_hierarchy', 'grid2mesh_edges', 'mesh2mesh_edges', 'mesh2grid_edges'} of one trajectory's
    initial frame (reference :27-60); tensors live on the device of `initial_positions` (CPU input is
    computed on the current CUDA device and returned on the CPU, as MultiScaleGraph does)."""
    config = MultiScaleConfig(num_scales=num_scales, window_size=window_size, radius_multiplier=radius_multiplier)
    return MultiScaleGraph(config).create_all_edges(initial_positions)


def _static_graphs(trajectories, device, num_scales, window_size, radius_multiplier, verbose=True):
    """One graph per trajectory from its frame 0 (reference :98-118, :168-188)."""
    if verbose:
        print(f"Building static multi-scale graphs for {len(trajectories)} trajectories...")
    graphs = {}
    for i, (positions, _, _) in enumerate(trajectories):
        p0 = torch.as_tensor(np.ascontiguousarray(positions[0], dtype=np.float32)).to(device)
        graphs[i] = build_static_multi_scale_graph(p0, num_scales, window_size, radius_multiplier)
    if verbose:
        print(f"Built {len(graphs)} static multi-scale graphs (scales {num_scales}, window {window_size})")
    return graphs


class MultiScaleTaylorImpactSamplesDataset(D.DeviceSamples):
    """One-step training samples + the static graph of their trajectory (reference :63-138).
    Items: {'input', 'output', 'meta', 'graph'}; sample indexing as sgnn_b200.data.DeviceSamples."""

    def __init__(self, data_path, input_length_sequence: int = 3, load_stress_stats: bool = True,
                 num_scales: int = 3, window_size: int = 3, radius_multiplier: float = 2.0, device="cuda"):
        trs = D.load_trajectories(data_path) if isinstance(data_path, (str, Path)) else list(data_path)
        super().__init__(trs, input_length_sequence, device)
        self._num_scales, self._window_size = num_scales, window_size
        self._static_graphs = _static_graphs(trs, self.device, num_scales, window_size, radius_multiplier)

    def __getitem__(self, idx: int) -> Dict:
        """The reference's un-collated sample (taylor_impact_data_loader.py:143-181 + :134-136)."""
        tr, t = self.locate(int(idx))
        n = self.pos[tr].shape[1]
        return {"input": {"positions": self.pos[tr][t - self.s:t].permute(1, 0, 2).contiguous(),
                          "particle_type": torch.full((n,), self.ptype[tr], dtype=torch.int64, device=self.device),
                          "n_particles_per_example": n},
                "output": {"next_position": self.pos[tr][t], "next_strain": self.stress[tr][t]},
                "meta": {"trajectory_idx": tr, "time_idx": t},
                "graph": self._static_graphs[tr]}

    def batch(self, indices) -> Dict:
        """Collated batch with the graph of the FIRST sample (multi_scale_collate_fn, :212-232)."""
        out = super().batch(indices)
        out["graph"] = self._static_graphs[self.locate(int(indices[0]))[0]]
        return out


def multi_scale_collate_fn(batch: List[Dict]) -> Dict:
    """Collate a list of samples (reference :212-232): tensors concatenated as the single-scale
    collate_fn does (taylor_impact_data_loader.py:243-284), graph taken from batch[0]."""
    dev = batch[0]["input"]["positions"].device
    out = {
        "input": {"positions": torch.cat([b["input"]["positions"] for b in batch]).contiguous(),
                  "particle_type": torch.cat([b["input"]["particle_type"] for b in batch]),
                  "n_particles_per_example": torch.tensor([int(b["input"]["n_particles_per_example"]) for b in batch],
                                                          dtype=torch.int64, device=dev)},
        "output": {"next_position": torch.cat([b["output"]["next_position"] for b in batch]).contiguous(),
                   "next_strain": torch.cat([b["output"]["next_strain"] for b in batch]).contiguous()},
        "meta": {"trajectory_idx": torch.tensor([b["meta"]["trajectory_idx"] for b in batch]),
                 "time_idx": torch.tensor([b["meta"]["time_idx"] for b in batch])},
    }
    if "graph" in batch[0]:
        out["graph"] = batch[0]["graph"]
    return out


class MultiScaleTaylorImpactTrajectoriesDataset(D.TrajectoriesOnDevice):
    """Whole trajectories + their static graph for rollout evaluation (reference :141-209)."""

    def __init__(self, data_path, load_stress_stats: bool = True, num_scales: int = 3, window_size: int = 3,
                 radius_multiplier: float = 2.0, device="cuda"):
        super().__init__(data_path, device)
        self._num_scales, self._window_size = num_scales, window_size
        self._static_graphs = _static_graphs(self.trs, self.device, num_scales, window_size, radius_multiplier)

    def __getitem__(self, idx: int) -> Dict:
        item = super().__getitem__(idx)
        item["graph"] = self._static_graphs[idx]
        return item

    def __iter__(self) -> Iterator[Dict]:
        return (self[i] for i in range(len(self)))


def get_multi_scale_data_loader_by_samples(path: str, input_length_sequence: int = 3, batch_size: int = 2,
                                           shuffle: bool = True, num_workers: int = 0, pin_memory: bool = True,
                                           load_stress_stats: bool = True, num_scales: int = 3, window_size: int = 3,
                                           radius_multiplier: float = 2.0, device="cuda", seed: int = 0,
                                           rank: int = 0, world: int = 1) -> D.DeviceSampleLoader:
    """Reference :235-278.  Iterating yields collated batches (dicts) with a 'graph' entry."""
    ds = MultiScaleTaylorImpactSamplesDataset(path, input_length_sequence, load_stress_stats, num_scales,
                                              window_size, radius_multiplier, device)
    return D.DeviceSampleLoader(ds, batch_size, shuffle, seed, rank, world)


def get_multi_scale_data_loader_by_trajectories(path: str, num_workers: int = 0, pin_memory: bool = True,
                                                load_stress_stats: bool = True, num_scales: int = 3,
                                                window_size: int = 3, radius_multiplier: float = 2.0,
                                                device="cuda") -> MultiScaleTaylorImpactTrajectoriesDataset:
    """Reference :281-314 (a DataLoader with batch_size=None there: iterating yields the items)."""
    return MultiScaleTaylorImpactTrajectoriesDataset(path, load_stress_stats, num_scales, window_size,
                                                     radius_multiplier, device)
