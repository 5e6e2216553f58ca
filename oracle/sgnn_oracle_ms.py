"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's multi-scale variant
(sgnn/multi_scale/*).  Same rules as oracle/sgnn_oracle.py: a checker, never a product path.
"""
from typing import Dict, Optional

import numpy as np
import torch

try:
    from .radius import radius_graph_np
    from .sgnn_oracle import (SD, decoder_postprocess, interaction_network,
                              inverse_decoder_postprocess, mlp, mlp_ln, time_diff)
except ImportError:  # pragma: no cover
    from radius import radius_graph_np
    from sgnn_oracle import (SD, decoder_postprocess, interaction_network,
                             inverse_decoder_postprocess, mlp, mlp_ln, time_diff)

GRID_SPACING = 0.5      # multi_scale_graph.py:34 (hard-coded)
MAX_NEIGHBORS = 24      # multi_scale_graph.py:36


def build_hierarchy(grid_positions: torch.Tensor, num_scales: int, window_size: int) -> Dict:
    """MultiScaleGraph.build_hierarchy/_sample_coarser_scale (multi_scale_graph.py:96-191):
    every window_size-th unique x and y coordinate, chained to global grid indices."""
    if num_scales < 2:
        raise ValueError(f"num_scales must be >= 2 (need grid + at least 1 mesh level), got {num_scales}")
    n = grid_positions.shape[0]
    hier = {0: {"sampling_indices": torch.arange(n, dtype=torch.long), "spacing": GRID_SPACING,
                "num_particles": n}}
    cur, spacing = grid_positions, GRID_SPACING
    for scale in range(1, num_scales):
        ux = torch.unique(cur[:, 0])  # sorted
        uy = torch.unique(cur[:, 1])
        mask = torch.isin(cur[:, 0], ux[::window_size]) & torch.isin(cur[:, 1], uy[::window_size])
        local = torch.where(mask)[0]
        glob = hier[scale - 1]["sampling_indices"][local]
        cur = cur[local]
        spacing = spacing * window_size
        hier[scale] = {"sampling_indices": glob, "spacing": spacing, "num_particles": cur.shape[0]}
    return hier


def _rg(pos: torch.Tensor, r: float) -> torch.Tensor:
    ei = radius_graph_np(pos.detach().to(torch.float32).numpy(), r, None, loop=True,
                         max_num_neighbors=MAX_NEIGHBORS)
    return torch.from_numpy(ei)


def create_all_edges(grid_positions: torch.Tensor, num_scales: int = 3, window_size: int = 3,
                     radius_multiplier: float = 2.0) -> Dict:
    """MultiScaleGraph.create_all_edges (multi_scale_graph.py:47-94, 193-281)."""
    hier = build_hierarchy(grid_positions, num_scales, window_size)
    mesh = hier[1]["sampling_indices"]
    grid_ei = _rg(grid_positions, radius_multiplier * GRID_SPACING)
    g2m = grid_ei[:, torch.isin(grid_ei[1], mesh)]
    m2g = grid_ei[:, torch.isin(grid_ei[0], mesh)]
    m2m = []
    for scale in range(1, num_scales):
        idx = hier[scale]["sampling_indices"]
        ei = _rg(grid_positions[idx], hier[scale]["spacing"] * radius_multiplier)
        if ei.shape[1] > 0:
            m2m.append(torch.stack([idx[ei[0]], idx[ei[1]]]))
    m2m = torch.cat(m2m, dim=1) if m2m else torch.empty((2, 0), dtype=torch.long)
    return {"graph_hierarchy": hier, "grid2mesh_edges": g2m, "mesh2mesh_edges": m2m,
            "mesh2grid_edges": m2g}


def multi_scale_features(position_sequence, graph: Dict, stats: Dict, num_scales: int,
                         radius_multiplier: float, particle_types=None,
                         type_embedding: Optional[torch.Tensor] = None):
    """MultiScaleSimulator._build_node_features/_build_edge_features
    (multi_scale_simulator.py:173-238): wall distance normalised by the grid radius;
    g2m/m2g displacements / grid radius, ALL m2m displacements / coarsest-scale radius."""
    n = position_sequence.shape[0]
    last = position_sequence[:, -1]
    vel = time_diff(position_sequence)
    vn = (vel - stats["velocity"]["mean"]) / stats["velocity"]["std"]
    rg = GRID_SPACING * radius_multiplier
    feats = [vn.reshape(n, -1), torch.clamp(last[:, 0:1] + 2.0, min=0.0, max=rg) / rg]
    if type_embedding is not None and type_embedding.shape[0] > 1:
        feats.append(type_embedding[particle_types])
    hier = graph["graph_hierarchy"]
    coarsest = num_scales - 1
    rc = hier[coarsest]["spacing"] * radius_multiplier if coarsest in hier else rg

    def ef(ei, r):
        if ei.shape[1] == 0:
            return torch.empty((0, 3), dtype=last.dtype)
        d = (last[ei[0]] - last[ei[1]]) / r
        return torch.cat([d, torch.norm(d, dim=-1, keepdim=True)], dim=-1)

    return torch.cat(feats, dim=-1), {"g2m": ef(graph["grid2mesh_edges"], rg),
                                      "m2m": ef(graph["mesh2mesh_edges"], rc),
                                      "m2g": ef(graph["mesh2grid_edges"], rg)}


def multi_scale_gnn(sd: SD, prefix: str, x, g2m_ei, g2m_ef, m2m_ei, m2m_ef, m2g_ei, m2g_ef,
                    return_latents: bool = False):
    """MultiScaleGNN.forward (multi_scale_gnn.py:277-325)."""
    h = mlp_ln(sd, prefix + "grid_node_encoder.", x)
    e_g2m = mlp_ln(sd, prefix + "g2m_edge_encoder.", g2m_ef)
    e_m2m = mlp_ln(sd, prefix + "m2m_edge_encoder.", m2m_ef)
    e_m2g = mlp_ln(sd, prefix + "m2g_edge_encoder.", m2g_ef)
    lat = [h]
    h, e_g2m = interaction_network(sd, prefix + "g2m_block.", h, g2m_ei, e_g2m)
    lat.append(h)
    layer = 0
    while f"{prefix}m2m_blocks.{layer}.edge_fn.0.NN-0.weight" in sd:
        h, e_m2m = interaction_network(sd, f"{prefix}m2m_blocks.{layer}.", h, m2m_ei, e_m2m)
        lat.append(h)
        layer += 1
    h, e_m2g = interaction_network(sd, prefix + "m2g_block.", h, m2g_ei, e_m2g)
    lat.append(h)
    out = mlp(sd, prefix + "prediction_head.", h)
    return (out, lat) if return_latents else out


def predict_positions(sd: SD, position_sequence, particle_types, graph: Dict, stats: Dict,
                      num_scales: int, radius_multiplier: float = 2.0, dim: int = 2):
    """MultiScaleSimulator.predict_positions (multi_scale_simulator.py:267-310)."""
    emb = sd.get("_particle_type_embedding.weight")
    nf, ef = multi_scale_features(position_sequence, graph, stats, num_scales, radius_multiplier,
                                  particle_types, emb)
    pred = multi_scale_gnn(sd, "_multi_scale_gnn.", nf, graph["grid2mesh_edges"], ef["g2m"],
                           graph["mesh2mesh_edges"], ef["m2m"], graph["mesh2grid_edges"], ef["m2g"])
    return decoder_postprocess(pred[:, :dim], position_sequence, stats), pred[:, -1]


def predict_accelerations(sd: SD, next_positions, noise, position_sequence, particle_types,
                          graph: Dict, stats: Dict, num_scales: int,
                          radius_multiplier: float = 2.0, dim: int = 2):
    """MultiScaleSimulator.predict_accelerations (multi_scale_simulator.py:312-345)."""
    noisy = position_sequence + noise
    emb = sd.get("_particle_type_embedding.weight")
    nf, ef = multi_scale_features(noisy, graph, stats, num_scales, radius_multiplier,
                                  particle_types, emb)
    pred = multi_scale_gnn(sd, "_multi_scale_gnn.", nf, graph["grid2mesh_edges"], ef["g2m"],
                           graph["mesh2mesh_edges"], ef["m2m"], graph["mesh2grid_edges"], ef["m2g"])
    target = inverse_decoder_postprocess(next_positions + noise[:, -1], noisy, stats)
    return pred[:, :dim], target, pred[:, -1]


def rollout(sd: SD, positions, particle_types, nsteps: int, graph: Dict, stats: Dict,
            input_sequence_length: int, num_scales: int, radius_multiplier: float = 2.0,
            dim: int = 2):
    """evaluate_multi_scale_rollout loop (multi_scale_evaluate.py:164-220), autoregressive."""
    S = input_sequence_length
    cur = positions[:, :S]
    preds, strains = [], []
    for _ in range(nsteps):
        nxt, st = predict_positions(sd, cur, particle_types, graph, stats, num_scales,
                                    radius_multiplier, dim)
        preds.append(nxt)
        strains.append(st)
        cur = torch.cat([cur[:, 1:], nxt[:, None, :]], dim=1)
    return torch.stack(preds), torch.stack(strains)
