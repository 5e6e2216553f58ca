"""MultiScaleSimulator with the reference's API (sgnn/multi_scale/multi_scale_simulator.py): static
multi-scale graph, per-step feature build, MultiScaleGNN, Euler integrator -- one forward is
3 edge-encoder launches + 1 node-encoder launch + 2 launches per block, no host synchronisation.

Observable differences: CUDA sm_100 tensors only (no CPU path); extra kwarg
`precision`; the static graph's edge lists are sorted by receiver ONCE in set_static_graph (stable,
so each receiver sums its messages in the caller's edge order) instead of being copied to the
device on every step (reference :157-159).
"""
from typing import Any, Dict, Optional, Tuple

import torch
import torch.nn as nn

from .. import _lib as L
from ..blob import params_version
from ..engine import Csr, EpdEngine
from ..single_scale.learned_simulator import example_ptr_of  # noqa: F401  (re-exported helper)
from .multi_scale_gnn import MultiScaleGNN
from .multi_scale_graph import MultiScaleConfig

_KEYS = ("grid2mesh_edges", "mesh2mesh_edges", "mesh2grid_edges")


class MultiScaleSimulator(nn.Module):
    """Reference class of the same name (multi_scale_simulator.py:20-388)."""

    def __init__(self, kinematic_dimensions: int, nnode_in: int, nedge_in: int, nedge_out: int,
                 latent_dim: int, nmessage_passing_steps: int, nmlp_layers: int,
                 normalization_stats: Dict, nparticle_types: int, particle_type_embedding_size: int,
                 num_scales: int = 3, window_size: int = 3, radius_multiplier: float = 2.0,
                 device: str = "cpu", precision: str = "tf32"):
        super().__init__()
        self._kinematic_dimensions = kinematic_dimensions
        self._normalization_stats = normalization_stats
        self._nparticle_types = nparticle_types
        self._num_scales = num_scales
        self._window_size = window_size
        self._device = device
        self._nnode_in, self._latent_dim = nnode_in, latent_dim
        self._particle_type_embedding = nn.Embedding(nparticle_types, particle_type_embedding_size)
        self._multi_scale_config = MultiScaleConfig(num_scales=num_scales, window_size=window_size,
                                                    radius_multiplier=radius_multiplier)
        self._multi_scale_gnn = MultiScaleGNN(
            nnode_in_features=nnode_in, nnode_out_features=kinematic_dimensions + 1,
            nedge_in_features=nedge_in, nedge_out_features=nedge_out, latent_dim=latent_dim,
            nmessage_passing_steps=nmessage_passing_steps, nmlp_layers=nmlp_layers, num_scales=num_scales)
        self._static_graph_data = None
        self.precision = precision
        self._graph_dev = None          # (device, [Csr x3], [perm x3]) of the current static graph
        self._engines = {}
        self._stats_dev = None

    def forward(self):
        pass

    # ---------------------------------------------------------------- static graph
    def set_static_graph(self, graph_data: Dict[str, Any]):
        """Store the static graph (reference :103-113); device CSRs are built lazily, once."""
        self._static_graph_data = graph_data
        self._graph_dev = None
        self._graph_token = getattr(self, "_graph_token", 0) + 1
        self.__dict__.pop("_rollout_plans", None)       # captured graphs point at the old CSRs

    def get_static_graph_data(self) -> Optional[Dict[str, Any]]:
        return self._static_graph_data

    def _validate_static_graph(self):
        if self._static_graph_data is None:
            raise ValueError("Static graph data not set. Call set_static_graph() first.")
        for key in ("graph_hierarchy",) + _KEYS:
            if key not in self._static_graph_data:
                raise ValueError(f"Missing required graph data key: {key}")

    def _radii(self) -> Tuple[float, float]:
        """(grid radius, coarsest-scale radius): g2m/m2g displacements are divided by the first,
        ALL m2m displacements by the second (reference :213-219,:234-236)."""
        cfg = self._multi_scale_config
        rg = cfg.grid_spacing * cfg.radius_multiplier
        hier = self._static_graph_data["graph_hierarchy"]
        coarsest = cfg.num_scales - 1
        rc = hier[coarsest]["spacing"] * cfg.radius_multiplier if coarsest in hier else rg
        return float(rg), float(rc)

    def _graph(self, n: int, device):
        if self._graph_dev is None or self._graph_dev[0] != (n, device):
            csrs, perms = [], []
            for key in _KEYS:
                ei = self._static_graph_data[key].to(device)
                csr, perm = Csr.from_edge_index(ei, n)
                csrs.append(csr)
                perms.append(perm)
            self._graph_dev = ((n, device), csrs, perms)
        return self._graph_dev[1], self._graph_dev[2]

    # ---------------------------------------------------------------- plumbing
    def _stats(self, device):
        if self._stats_dev is None or self._stats_dev[0] != device:
            f = lambda t: torch.as_tensor(t).detach().to(device=device, dtype=torch.float32).contiguous()  # noqa: E731
            self._stats_dev = (device, {k: {kk: f(vv) for kk, vv in v.items()}
                                        for k, v in self._normalization_stats.items()})
        return self._stats_dev[1]

    def _engine(self, n: int, device, csrs):
        caps = tuple(c.max_edges for c in csrs)
        key = (n, device, self.precision, caps)
        ver = params_version(self.parameters())
        hit = self._engines.get(key)
        if hit is None or hit[0] != ver:
            g = self._multi_scale_gnn
            blocks = [(g.g2m_block.edge_fn, g.g2m_block.node_fn, 0, 0)]
            blocks += [(b.edge_fn, b.node_fn, 1, l) for l, b in enumerate(g.m2m_blocks)]
            blocks += [(g.m2g_block.edge_fn, g.m2g_block.node_fn, 2, 0)]
            eng = EpdEngine(g.grid_node_encoder, [g.g2m_edge_encoder, g.m2m_edge_encoder, g.m2g_edge_encoder],
                            blocks, g.prediction_head, k_node_in=self._nnode_in,
                            dim=self._kinematic_dimensions, latent=self._latent_dim, n_nodes=n,
                            edge_caps=list(caps), precision=self.precision, device=device,
                            has_embedding=self._nparticle_types > 1)
            hit = (ver, eng)
            self._engines = {key: hit}
        return hit[1]

    def _prepare(self, n: int, device):
        """Build / refresh the static-graph CSRs and kernel plans outside a CUDA-graph capture."""
        self._validate_static_graph()
        csrs, _ = self._graph(n, device)
        self._engine(n, device, csrs)

    @staticmethod
    def _window(position_sequence):
        seq = position_sequence
        if seq.dtype != torch.float32:
            seq = seq.float()
        if seq.stride(2) != 1 or seq.stride(1) != seq.shape[2]:
            seq = seq.contiguous()
        return seq

    def _step(self, position_sequence, nparticles_per_example, particle_types, next_pos=None,
              strain=None, feat_out=None, edge_feat_out=None, example_ptr=None):
        """Whole forward on the current stream.  Returns (engine, csrs)."""
        self._validate_static_graph()
        L.require_cuda(position_sequence, "position_sequence")
        seq = position_sequence
        n, s, dim = seq.shape
        dev = seq.device
        csrs, _ = self._graph(n, dev)
        eng = self._engine(n, dev, csrs)
        st = self._stats(dev)
        rg, rc = self._radii()
        use_emb = self._nparticle_types > 1
        if use_emb:      # the kernel indexes the embedding table with these ids (clamped to the table there)
            L.require_cuda(particle_types, "particle_types")
            if particle_types.dtype != torch.int64 or not particle_types.is_contiguous():
                particle_types = particle_types.to(torch.int64).contiguous()
            if particle_types.numel() != n:
                raise ValueError(f"particle_types has {particle_types.numel()} entries for {n} particles")
        stride = seq.stride(0)
        eng.forward(pos_seq=seq, pos_seq_stride=stride, seq_len=s, csrs=csrs, radii=[rg, rc, rg],
                    vel_mean=st["velocity"]["mean"], vel_std=st["velocity"]["std"], wall_max=rg,
                    wall_div=rg, types=particle_types if use_emb else None,
                    emb=self._particle_type_embedding.weight.detach() if use_emb else None,
                    acc_mean=st["acceleration"]["mean"], acc_std=st["acceleration"]["std"],
                    next_pos=next_pos, next_pos_stride=0 if next_pos is None else next_pos.stride(0),
                    strain_out=strain, feat_out=feat_out, edge_feat_out=edge_feat_out)
        return eng, csrs

    # ---------------------------------------------------------------- reference API
    def _encoder_preprocessor(self, position_sequence, nparticles_per_example, particle_types):
        """(node_features, edge_indices, edge_features) as the reference builds them (:125-171);
        edge features are returned in the caller's edge order."""
        self._validate_static_graph()
        L.require_cuda(position_sequence, "position_sequence")
        seq = self._window(position_sequence.detach())
        n, dev = seq.shape[0], seq.device
        csrs, perms = self._graph(n, dev)
        k = self._kinematic_dimensions + 1
        nf = torch.empty(n, self._nnode_in, dtype=torch.float32, device=dev)
        ef_csr = [torch.empty(c.max_edges, k, dtype=torch.float32, device=dev) for c in csrs]
        self._step(seq, nparticles_per_example, particle_types, feat_out=nf, edge_feat_out=ef_csr)
        edge_indices, edge_features = {}, {}
        for name, key, perm, ef in zip(("g2m", "m2m", "m2g"), _KEYS, perms, ef_csr):
            ei = self._static_graph_data[key].to(dev)
            e = ei.shape[1]
            out = torch.empty(e, k, dtype=torch.float32, device=dev)
            out[perm.long()] = ef[:e]
            edge_indices[name], edge_features[name] = ei, out
        return nf, edge_indices, edge_features

    def predict_positions(self, current_positions, nparticles_per_example, particle_types):
        """Next positions and the auxiliary strain channel (reference :267-310)."""
        L.require_cuda(current_positions, "current_positions")
        seq = self._window(current_positions.detach())
        n, _, dim = seq.shape
        next_pos = torch.empty(n, dim, dtype=torch.float32, device=seq.device)
        strain = torch.empty(n, dtype=torch.float32, device=seq.device)
        self._step(seq, nparticles_per_example, particle_types, next_pos=next_pos, strain=strain)
        return next_pos, strain

    def predict_accelerations(self, next_positions, position_sequence_noise, position_sequence,
                              nparticles_per_example, particle_types):
        """Training forward with autograd (reference :312-345); see sgnn_b200/train_ops.py."""
        from .. import train_ops as T
        self._validate_static_graph()
        L.require_cuda(position_sequence, "position_sequence")
        noisy = self._window((position_sequence + position_sequence_noise).detach())
        n, s, dim = noisy.shape
        dev = noisy.device
        st = self._stats(dev)
        rg, rc = self._radii()
        csrs, _ = self._graph(n, dev)
        tg = self.__dict__.get("_train_graphs")
        if tg is None or tg[0] is not csrs:
            tg = self.__dict__["_train_graphs"] = (csrs, [T.TrainGraph.from_csr(c) for c in csrs])
        graphs = tg[1]
        extra = self._particle_type_embedding(particle_types) if self._nparticle_types > 1 else None
        nf = T.node_features(noisy, st["velocity"]["mean"], st["velocity"]["std"], rg, rg, extra)
        last = noisy[0, s - 1]
        efs = [T.edge_features(last, noisy.stride(0), dim, r, g) for r, g in zip((rg, rc, rg), graphs)]
        g = self._multi_scale_gnn
        blocks = [(g.g2m_block.edge_fn, g.g2m_block.node_fn, 0)]
        blocks += [(b.edge_fn, b.node_fn, 1) for b in g.m2m_blocks]
        blocks += [(g.m2g_block.edge_fn, g.m2g_block.node_fn, 2)]
        pred = T.epd_forward(g.grid_node_encoder, [g.g2m_edge_encoder, g.m2m_edge_encoder, g.m2g_edge_encoder],
                             blocks, g.prediction_head, nf, efs, graphs)
        next_adj = next_positions + position_sequence_noise[:, -1]
        target = self._inverse_decoder_postprocessor(next_adj, noisy)
        return pred[:, :dim], target, pred[:, -1]

    def _decoder_postprocessor(self, normalized_acceleration, position_sequence):
        st = self._stats(normalized_acceleration.device)["acceleration"]
        acceleration = normalized_acceleration * st["std"] + st["mean"]
        last = position_sequence[:, -1]
        return last + ((last - position_sequence[:, -2]) + acceleration)

    def _inverse_decoder_postprocessor(self, next_position, position_sequence):
        st = self._stats(next_position.device)["acceleration"]
        prev = position_sequence[:, -1]
        acceleration = (next_position - prev) - (prev - position_sequence[:, -2])
        return (acceleration - st["mean"]) / st["std"]

    def _time_diff(self, position_sequence: torch.Tensor) -> torch.Tensor:
        return (position_sequence[:, 1:] - position_sequence[:, :-1]).contiguous()

    def save(self, path: str = "multi_scale_model.pt"):
        torch.save(self.state_dict(), path)

    def load(self, path: str):
        self.load_state_dict(torch.load(path, map_location=torch.device("cpu")))
