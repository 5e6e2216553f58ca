"""Multi-scale Encode-Process-Decode network with the reference's module API and checkpoint keys
(sgnn/multi_scale/multi_scale_gnn.py), executed by the sm_100a kernels of libsgnn_b200.

The reference's "pooling/unpooling" is three InteractionNetwork-style blocks (G2M, L x M2M, M2G)
over ONE node array [N_grid, H] (multi_scale_gnn.py:301-325); the three block classes differ only
in name.  As under PyG's propagate, every block returns `edge_features + edge_features`
(update() hands back its input, :84-110).  No PyTorch/CPU fallback.

The fused per-step path used by MultiScaleSimulator lives in sgnn_b200/engine.py; forward() here
serves callers that drive the network with explicit feature tensors and arbitrary (unsorted,
duplicated) edge lists (sgnn/multi_scale/test/test_multi_scale_gnn.py:16-71).
"""
import ctypes as C
from typing import Any, Dict

import torch
import torch.nn as nn

from .. import _lib as L
from ..engine import Csr, EdgeEncoderPlan, NodePlan, TILE
from ..single_scale.graph_network import InteractionNetwork, _PlanCache, _f32c, _rows_forward, build_mlp

__all__ = ["build_mlp", "G2MBlock", "M2MBlock", "M2GBlock", "MultiScaleGNN"]


class _Block(InteractionNetwork):
    """Shared body of G2MBlock / M2MBlock / M2GBlock (reference :66-205): edge MLP on
    [x_target | x_source | e], sum at the target, node MLP on [aggregate | x] over ALL nodes,
    residual.  Hidden widths are `latent_dim`, repeated `nmlp_layers` times."""

    def __init__(self, nnode_in: int, nnode_out: int, nedge_in: int, nedge_out: int, nmlp_layers: int,
                 latent_dim: int):
        super().__init__(nnode_in=nnode_in, nnode_out=nnode_out, nedge_in=nedge_in, nedge_out=nedge_out,
                         nmlp_layers=nmlp_layers, mlp_hidden_dim=latent_dim)


class G2MBlock(_Block):
    """Grid -> mesh block (reference G2MBlock, multi_scale_gnn.py:66-110)."""


class M2MBlock(_Block):
    """Mesh -> mesh processor block (reference M2MBlock, :113-157)."""


class M2GBlock(_Block):
    """Mesh -> grid block (reference M2GBlock, :160-205)."""


def _encode_edges(cache: _PlanCache, owner: nn.Module, seq: nn.Sequential, ef: torch.Tensor):
    """Edge-encoder MLP + LayerNorm on explicit [E, k] features -> [E, out] rows."""
    ef = _f32c(ef, "edge_features")
    dev, e = ef.device, ef.shape[0]
    plan = cache.get(owner, lambda tc: EdgeEncoderPlan(seq, ef.shape[1], tc, dev))
    out = torch.empty(e, plan.out, dtype=torch.float32, device=dev)
    if e == 0:
        return out
    csr = Csr(1, e, dev)                   # capacity only: explicit features, no gather
    tiles = torch.empty(csr.max_tiles * TILE * plan.out, dtype=torch.float32, device=dev)
    plan.launch(csr, tiles, feat=ef, n_edges=e)
    L.check(L.lib().sgnn_unpack_tiles_to_rows(L.ptr(tiles), None, e, plan.out, C.c_float(1.0),
                                              L.ptr(out), C.c_long(plan.out), L.stream_ptr()))
    return out


class MultiScaleGNN(nn.Module):
    """Grid encoder + 3 edge encoders -> G2M -> L x M2M -> M2G -> prediction head
    (reference MultiScaleGNN, multi_scale_gnn.py:208-325)."""

    def __init__(self, nnode_in_features: int, nnode_out_features: int, nedge_in_features: int,
                 nedge_out_features: int, latent_dim: int, nmessage_passing_steps: int,
                 nmlp_layers: int, num_scales: int, share_weights_across_scales: bool = False):
        super().__init__()
        self.num_scales = num_scales
        self.latent_dim = latent_dim
        self.nmessage_passing_steps = nmessage_passing_steps
        hidden = [latent_dim for _ in range(nmlp_layers)]
        self.grid_node_encoder = nn.Sequential(build_mlp(nnode_in_features, hidden, latent_dim),
                                               nn.LayerNorm(latent_dim))
        self.g2m_edge_encoder = nn.Sequential(build_mlp(nedge_in_features, hidden, nedge_out_features),
                                              nn.LayerNorm(nedge_out_features))
        self.m2m_edge_encoder = nn.Sequential(build_mlp(nedge_in_features, hidden, nedge_out_features),
                                              nn.LayerNorm(nedge_out_features))
        self.m2g_edge_encoder = nn.Sequential(build_mlp(nedge_in_features, hidden, nedge_out_features),
                                              nn.LayerNorm(nedge_out_features))
        self.g2m_block = G2MBlock(latent_dim, latent_dim, nedge_out_features, nedge_out_features,
                                  nmlp_layers, latent_dim)
        self.m2m_blocks = nn.ModuleList([
            M2MBlock(latent_dim, latent_dim, nedge_out_features, nedge_out_features, nmlp_layers, latent_dim)
            for _ in range(nmessage_passing_steps)])
        self.m2g_block = M2GBlock(latent_dim, latent_dim, nedge_out_features, nedge_out_features,
                                  nmlp_layers, latent_dim)
        self.prediction_head = build_mlp(latent_dim, hidden, nnode_out_features)
        # plan caches (plain attributes: no parameters, no state_dict entries)
        self._caches = {k: _PlanCache() for k in ("node", "g2m", "m2m", "m2g", "head")}

    def forward(self, x: torch.Tensor, g2m_edge_index: torch.Tensor, g2m_edge_features: torch.Tensor,
                m2m_edge_index: torch.Tensor, m2m_edge_features: torch.Tensor,
                m2g_edge_index: torch.Tensor, m2g_edge_features: torch.Tensor,
                graph_hierarchy: Dict[int, Dict[str, Any]] = None) -> torch.Tensor:
        x = _f32c(x, "x")
        dev = x.device
        node_plan = self._caches["node"].get(
            self.grid_node_encoder, lambda tc: NodePlan(L.PRO_ROWS, x.shape[1], self.grid_node_encoder, tc, dev))
        grid = _rows_forward(node_plan, x, node_plan.p.mlp.out)
        e_g2m = _encode_edges(self._caches["g2m"], self.g2m_edge_encoder, self.g2m_edge_encoder, g2m_edge_features)
        e_m2m = _encode_edges(self._caches["m2m"], self.m2m_edge_encoder, self.m2m_edge_encoder, m2m_edge_features)
        e_m2g = _encode_edges(self._caches["m2g"], self.m2g_edge_encoder, self.m2g_edge_encoder, m2g_edge_features)
        grid, e_g2m = self.g2m_block(grid, g2m_edge_index, e_g2m)
        for blk in self.m2m_blocks:
            grid, e_m2m = blk(grid, m2m_edge_index, e_m2m)
        grid, e_m2g = self.m2g_block(grid, m2g_edge_index, e_m2g)
        head = self._caches["head"].get(
            self.prediction_head, lambda tc: NodePlan(L.PRO_ROWS, grid.shape[1], None, tc, dev,
                                                      decoder=self.prediction_head))
        return _rows_forward(head, grid, head.p.dec.out)
