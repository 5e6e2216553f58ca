"""Static multi-scale graph hierarchy with the reference's API
(sgnn/multi_scale/multi_scale_graph.py): lattice subsampling by every `window_size`-th unique
coordinate, grid radius graph filtered into grid2mesh / mesh2grid edges, one radius graph per mesh
scale mapped to global grid ids and concatenated (duplicates across scales are kept, :75-85).

All radius graphs run on the sm_100a neighbour-search kernel (sgnn_radius_graph_csr, K = 24,
loop=True); the index bookkeeping (unique / isin / where) is torch on the same device.  Inputs on
the CPU are moved to the current CUDA device and the results are returned on the input's device,
so the reference's data loaders (static_graph_data_loader.py:101-113) can call it unchanged; there
is no CPU compute path.
"""
from typing import Any, Dict, Tuple

import torch

from .. import _lib as L
from ..engine import RadiusGraph

__all__ = ["MultiScaleConfig", "MultiScaleGraph"]


class MultiScaleConfig:
    """Configuration of the multi-scale mesh (reference MultiScaleConfig, :14-36)."""

    def __init__(self, num_scales: int = 3, window_size: int = 3, radius_multiplier: float = 2.0):
        if num_scales < 2:
            raise ValueError(f"num_scales must be >= 2 (need grid + at least 1 mesh level), got {num_scales}")
        self.num_scales = num_scales
        self.window_size = window_size
        self.grid_spacing = 0.5             # hard-coded in the reference (:34)
        self.radius_multiplier = radius_multiplier
        self.max_neighbors = 24             # (:36)


def _radius_graph(pos: torch.Tensor, r: float, max_neighbors: int) -> torch.Tensor:
    """[2, E] int64 (source; target), target-sorted: radius_graph(pos, r, loop=True, K)."""
    n, dim = pos.shape
    if n == 0:
        raise ValueError("radius graph of an empty particle set")
    pos = pos.detach().float().contiguous()
    rg = RadiusGraph(n, max_neighbors, True, pos.device)
    ptr = torch.tensor([0, n], dtype=torch.int32, device=pos.device)
    return rg.build(pos, dim, dim, ptr, float(r)).edge_index()


class MultiScaleGraph:
    """Reference MultiScaleGraph (:39-281)."""

    def __init__(self, config: MultiScaleConfig):
        self.config = config
        self.grid_positions = None
        self.graph_hierarchy = {}

    @staticmethod
    def _to_cuda(t: torch.Tensor) -> torch.Tensor:
        if t.is_cuda:
            return t
        if not torch.cuda.is_available():
            raise L.SgnnError("MultiScaleGraph needs a CUDA device: sgnn_b200 has no CPU path")
        return t.cuda()

    def create_all_edges(self, grid_positions: torch.Tensor) -> Dict[str, Any]:
        """All edge types in global grid indices (reference :47-94)."""
        out_dev = grid_positions.device
        if not self.graph_hierarchy:
            self.build_hierarchy(grid_positions)
        g2m, m2g = self._create_grid_mesh_connectivity(grid_positions)
        m2m = []
        for scale in range(1, self.config.num_scales):
            ei = self._create_mesh2mesh_edges(scale)
            if ei.shape[1] > 0:
                m2m.append(ei)
        m2m = torch.cat(m2m, dim=1) if m2m else torch.empty((2, 0), dtype=torch.long, device=out_dev)
        return {"graph_hierarchy": self.graph_hierarchy, "grid2mesh_edges": g2m.to(out_dev),
                "mesh2mesh_edges": m2m.to(out_dev), "mesh2grid_edges": m2g.to(out_dev)}

    def build_hierarchy(self, grid_positions: torch.Tensor) -> Dict[str, Any]:
        """Scale 0 = all grid particles, scale s = every window_size-th unique x and y coordinate of
        scale s-1, as global grid indices (reference :96-137)."""
        self.grid_positions = grid_positions
        dev = grid_positions.device
        self.graph_hierarchy[0] = {
            "sampling_indices": torch.arange(len(grid_positions), dtype=torch.long, device=dev),
            "spacing": self.config.grid_spacing, "num_particles": len(grid_positions)}
        cur, spacing = grid_positions, self.config.grid_spacing
        for scale in range(1, self.config.num_scales):
            cur, spacing, glob = self._sample_coarser_scale(cur, spacing, scale)
            self.graph_hierarchy[scale] = {"sampling_indices": glob, "spacing": spacing,
                                           "num_particles": len(cur)}
        return self.graph_hierarchy

    def _sample_coarser_scale(self, current_positions: torch.Tensor, current_spacing: float,
                              scale: int) -> Tuple[torch.Tensor, float, torch.Tensor]:
        """Reference :139-191 (requires exact lattice coordinates in frame 0)."""
        w = self.config.window_size
        x, y = current_positions[:, 0], current_positions[:, 1]
        ux, uy = torch.unique(x), torch.unique(y)        # sorted
        mask = torch.isin(x, ux[::w]) & torch.isin(y, uy[::w])
        local = torch.where(mask)[0]
        parent = self.graph_hierarchy[scale - 1]["sampling_indices"]
        return current_positions[local], current_spacing * w, parent[local]

    def _create_grid_mesh_connectivity(self, grid_positions: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
        """One grid radius graph (r = multiplier * 0.5), filtered on target in mesh (g2m) and on
        source in mesh (m2g) (reference :193-243)."""
        if 1 not in self.graph_hierarchy:
            raise ValueError("First mesh level 1 not found")
        pos = self._to_cuda(grid_positions)
        mesh = self.graph_hierarchy[1]["sampling_indices"].to(pos.device)
        ei = _radius_graph(pos, self.config.radius_multiplier * self.config.grid_spacing,
                           self.config.max_neighbors)
        return ei[:, torch.isin(ei[1], mesh)], ei[:, torch.isin(ei[0], mesh)]

    def _create_mesh2mesh_edges(self, scale: int) -> torch.Tensor:
        """Radius graph inside one mesh scale (r = spacing_s * multiplier), global ids (:245-281)."""
        if scale not in self.graph_hierarchy:
            raise ValueError(f"Scale {scale} not found")
        data = self.graph_hierarchy[scale]
        pos_all = self._to_cuda(self.grid_positions)
        idx = data["sampling_indices"].to(pos_all.device)
        ei = _radius_graph(pos_all[idx], data["spacing"] * self.config.radius_multiplier,
                           self.config.max_neighbors)
        return torch.stack([idx[ei[0]], idx[ei[1]]])
