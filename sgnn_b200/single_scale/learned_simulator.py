"""LearnedSimulator with the reference's API (sgnn/single_scale/learned_simulator.py), running
the whole step -- neighbour search, feature build, Encode-Process-Decode, Euler integrator --
as ~17 kernels of libsgnn_b200 with no host synchronisation.

Differences a caller can observe (all documented in DESIGN.md):
  * tensors must live on a CUDA (sm_100) device; there is no CPU path
  * `precision="tf32"` (default; tensor cores) or "fp32" (FFMA, tight parity) -- extra kwarg
  * the particle-count mismatch warning of the reference (:97-101) is not evaluated: it would
    need a device->host sync on every step
"""
from typing import Dict

import torch
import torch.nn as nn

from .. import _lib as L
from ..blob import params_version
from ..engine import EpdEngine, RadiusGraph
from . import graph_network

MAX_NUM_NEIGHBORS = 20      # hard-coded in the reference (learned_simulator.py:116-117)


def example_ptr_of(nparticles_per_example, device) -> torch.Tensor:
    """[B+1] int32 example boundaries on `device`, without forcing a sync for CUDA inputs.
    Accepts what the reference tolerates (:89-94; evaluate.py:121): a [B] tensor, a 0-d tensor,
    a list of ints, or a list holding 0-d tensors."""
    npe = nparticles_per_example
    if not isinstance(npe, torch.Tensor):
        if isinstance(npe, (list, tuple)) and any(isinstance(v, torch.Tensor) for v in npe):
            npe = torch.stack([torch.as_tensor(v).reshape(()) for v in npe])
        else:
            npe = torch.as_tensor(npe)
    npe = npe.flatten().to(device=device, dtype=torch.int32)
    ptr = torch.zeros(npe.numel() + 1, dtype=torch.int32, device=device)
    ptr[1:] = torch.cumsum(npe, 0)
    return ptr


class LearnedSimulator(nn.Module):
    """Learned simulator (https://arxiv.org/abs/2002.09405) -- reference class of the same name."""

    def __init__(self, particle_dimensions: int, nnode_in: int, nedge_in: int, latent_dim: int,
                 nmessage_passing_steps: int, nmlp_layers: int, mlp_hidden_dim: int,
                 connectivity_radius: float, normalization_stats: Dict, nparticle_types: int,
                 particle_type_embedding_size, device="cpu", precision: str = "tf32"):
        super().__init__()
        self._connectivity_radius = connectivity_radius
        self._normalization_stats = normalization_stats
        self._nparticle_types = nparticle_types
        self._particle_dimensions = particle_dimensions
        self._nnode_in, self._latent_dim = nnode_in, latent_dim
        self._particle_type_embedding = nn.Embedding(nparticle_types, particle_type_embedding_size)
        self._encode_process_decode = graph_network.EncodeProcessDecode(
            nnode_in_features=nnode_in,
            nnode_out_features=particle_dimensions + 1,   # accelerations + one auxiliary channel
            nedge_in_features=nedge_in,
            latent_dim=latent_dim,
            nmessage_passing_steps=nmessage_passing_steps,
            nmlp_layers=nmlp_layers,
            mlp_hidden_dim=mlp_hidden_dim)
        self._device = device
        self.precision = precision
        self._engines = {}
        self._stats_dev = None

    def forward(self):
        pass

    # ---------------------------------------------------------------- plumbing
    def _stats(self, device):
        if self._stats_dev is None or self._stats_dev[0] != device:
            s = self._normalization_stats
            f = lambda t: torch.as_tensor(t).detach().to(device=device, dtype=torch.float32).contiguous()  # noqa: E731
            self._stats_dev = (device, {k: {kk: f(vv) for kk, vv in v.items()} for k, v in s.items()})
        return self._stats_dev[1]

    def _engine(self, n: int, device, capacity: int = 0):
        """(EpdEngine, RadiusGraph) able to run n particles; rebuilt when parameters change.  An engine
        sized for somewhat more particles is reused (run with n_active = n): building one packs every
        weight image and allocates all latent buffers, which the sharded rollout cannot afford at every
        ghost refresh (_reserve)."""
        ver = params_version(self.parameters())
        key = (n, device, self.precision)
        hit = self._engines.get(key)
        if hit is None or hit[0] != ver:
            for (cap, d, pr), h in self._engines.items():
                if d == device and pr == self.precision and h[0] == ver and n <= cap <= n + max(n // 4, 4096):
                    return h[1], h[2]
        if hit is None or hit[0] != ver:
            size = max(n, capacity)
            key = (size, device, self.precision)
            old = self._engines.get(key)
            epd = self._encode_process_decode
            blocks = [(g.edge_fn, g.node_fn, 0, l) for l, g in enumerate(epd._processor.gnn_stacks)]
            rg = old[2] if old is not None else RadiusGraph(size, MAX_NUM_NEIGHBORS, True, device)
            eng = EpdEngine(epd._encoder.node_fn, [epd._encoder.edge_fn], blocks, epd._decoder.node_fn,
                            k_node_in=self._nnode_in, dim=self._particle_dimensions,
                            latent=self._latent_dim, n_nodes=size, edge_caps=[rg.csr.max_edges],
                            precision=self.precision, device=device,
                            has_embedding=self._nparticle_types > 1)
            hit = (ver, eng, rg)
            self._engines.pop(key, None)
            while len(self._engines) >= 2:                 # bounded cache: shapes change at halo refreshes
                self._engines.pop(next(iter(self._engines)))
            self._engines[key] = hit
        return hit[1], hit[2]

    def _reserve(self, capacity: int, device):
        """Kernel plans and buffers for up to `capacity` particles (reused for any smaller system)."""
        self._stats(device)
        self._engine(capacity, device, capacity)

    def _prepare(self, n: int, device):
        """Build / refresh kernel plans outside a CUDA-graph capture."""
        self._stats(device)
        self._engine(n, device)

    def _check_positions(self, position_sequence):
        if len(position_sequence.shape) != 3:
            raise ValueError(f"Expected position_sequence to have 3 dimensions, got {len(position_sequence.shape)}")
        if position_sequence.shape[1] < 2:
            raise ValueError(f"Expected at least 2 timesteps, got {position_sequence.shape[1]}")
        L.require_cuda(position_sequence, "position_sequence")

    def _step(self, position_sequence, nparticles_per_example, particle_types, next_pos=None,
              strain=None, feat_out=None, edge_feat_out=None, example_ptr=None):
        """Neighbour search + full network on the current stream.  Returns (engine, csr)."""
        seq = position_sequence
        n, s, dim = seq.shape
        dev = seq.device
        eng, rg = self._engine(n, dev)
        st = self._stats(dev)
        if example_ptr is None:
            example_ptr = example_ptr_of(nparticles_per_example, dev)
        stride = seq.stride(0)
        csr = rg.build(seq[0, s - 1], stride, dim, example_ptr, float(self._connectivity_radius), n=n)
        use_emb = self._nparticle_types > 1
        if use_emb:
            # the kernel indexes the embedding table with these ids: int64, on the device, contiguous.  Ids
            # outside [0, ntypes) are clamped by the kernel (nn.Embedding would raise; a fault on the device
            # cannot be raised as an exception without a host sync)
            L.require_cuda(particle_types, "particle_types")
            if particle_types.dtype != torch.int64 or not particle_types.is_contiguous():
                particle_types = particle_types.to(torch.int64).contiguous()
            if particle_types.numel() != n:
                raise ValueError(f"particle_types has {particle_types.numel()} entries for {n} particles")
        eng.forward(pos_seq=seq, pos_seq_stride=stride, seq_len=s, csrs=[csr], n_active=n,
                    radii=[float(self._connectivity_radius)], vel_mean=st["velocity"]["mean"],
                    vel_std=st["velocity"]["std"], wall_max=float(self._connectivity_radius),
                    wall_div=1.0, types=particle_types if use_emb else None,
                    emb=self._particle_type_embedding.weight.detach() if use_emb else None,
                    acc_mean=st["acceleration"]["mean"], acc_std=st["acceleration"]["std"],
                    next_pos=next_pos, next_pos_stride=0 if next_pos is None else next_pos.stride(0),
                    strain_out=strain, feat_out=feat_out,
                    edge_feat_out=None if edge_feat_out is None else [edge_feat_out])
        return eng, csr

    @staticmethod
    def _window(position_sequence):
        """fp32 view with unit stride over [S, dim] per particle (particle stride may be larger)."""
        seq = position_sequence
        if seq.dtype != torch.float32:
            seq = seq.float()
        if seq.stride(2) != 1 or seq.stride(1) != seq.shape[2]:
            seq = seq.contiguous()
        return seq

    # ---------------------------------------------------------------- reference API
    def _compute_graph_connectivity(self, positions, nparticles_per_example, radius: float,
                                    add_self_edges: bool = True):
        """Edges to all particles within `radius`, at most 20 per receiver, same example only
        (reference :70-124).  Returns (edge_index[0], edge_index[1]) = (sources, targets) of
        radius_graph, exactly as the reference returns them under the names receivers/senders."""
        if len(positions.shape) != 2:
            raise ValueError(f"Expected 2D positions tensor, got shape {positions.shape}")
        L.require_cuda(positions, "positions")
        pos = positions.detach().float().contiguous()
        n, dim = pos.shape
        rg = RadiusGraph(n, MAX_NUM_NEIGHBORS, add_self_edges, pos.device)
        csr = rg.build(pos, dim, dim, example_ptr_of(nparticles_per_example, pos.device), float(radius))
        ei = csr.edge_index()
        return ei[0], ei[1]

    def _encoder_preprocessor(self, position_sequence, nparticles_per_example, particle_types):
        """(node_features, edge_index, edge_features) as the reference builds them (:231-316)."""
        self._check_positions(position_sequence)
        seq = self._window(position_sequence.detach())
        n = seq.shape[0]
        eng, rg = self._engine(n, seq.device)
        nf = torch.empty(n, self._nnode_in, dtype=torch.float32, device=seq.device)
        ef = torch.empty(rg.csr.max_edges, self._particle_dimensions + 1, dtype=torch.float32, device=seq.device)
        _, csr = self._step(seq, nparticles_per_example, particle_types, feat_out=nf, edge_feat_out=ef)
        ei = csr.edge_index()
        return nf, ei, ef[:ei.shape[1]]

    def predict_positions(self, current_positions, nparticles_per_example, particle_types):
        """Next positions and auxiliary strain channel (reference :413-438)."""
        self._check_positions(current_positions)
        seq = self._window(current_positions.detach())
        n, _, dim = seq.shape
        next_pos = torch.empty(n, dim, dtype=torch.float32, device=seq.device)
        strain = torch.empty(n, dtype=torch.float32, device=seq.device)
        self._step(seq, nparticles_per_example, particle_types, next_pos=next_pos, strain=strain)
        return next_pos, strain

    def predict_accelerations(self, next_positions, position_sequence_noise, position_sequence,
                              nparticles_per_example, particle_types):
        """Training forward (reference :440-491): noisy inputs -> graph on the noisy last frame
        -> network with autograd -> (predicted normalised acceleration, target, predicted strain).
        Graph and features come from the sm_100a kernels, the message-passing glue and its
        backward from csrc/graph_ops.cu, GEMM/LayerNorm from cuBLAS/ATen (sgnn_b200/train_ops.py)."""
        from .. import train_ops as T
        self._check_positions(position_sequence)
        noisy = self._window((position_sequence + position_sequence_noise).detach())
        n, s, dim = noisy.shape
        dev = noisy.device
        st = self._stats(dev)
        radius = float(self._connectivity_radius)
        rg = self.__dict__.get("_train_rg")
        if rg is None or rg.n != n or rg.csr.rowptr.device != dev:
            rg = self.__dict__["_train_rg"] = RadiusGraph(n, MAX_NUM_NEIGHBORS, True, dev)
        csr = rg.build(noisy[0, s - 1], noisy.stride(0), dim, example_ptr_of(nparticles_per_example, dev), radius)
        graph = T.TrainGraph.from_csr(csr, copy=torch.is_grad_enabled())     # two forwards before one backward stay correct
        extra = self._particle_type_embedding(particle_types) if self._nparticle_types > 1 else None
        nf = T.node_features(noisy, st["velocity"]["mean"], st["velocity"]["std"], radius, 1.0, extra)
        ef = T.edge_features(noisy[0, s - 1], noisy.stride(0), dim, radius, graph)
        epd = self._encode_process_decode
        blocks = [(g.edge_fn, g.node_fn, 0) for g in epd._processor.gnn_stacks]
        pred = T.epd_forward(epd._encoder.node_fn, [epd._encoder.edge_fn], blocks, epd._decoder.node_fn,
                             nf, [ef], [graph])
        next_adj = next_positions + position_sequence_noise[:, -1]
        target = self._inverse_decoder_postprocessor(next_adj, noisy)
        return pred[:, :dim], target, pred[:, -1]

    def _decoder_postprocessor(self, normalized_acceleration, position_sequence):
        """Euler integrator on explicit tensors (reference :381-411); the fused step applies the
        same arithmetic inside the decoder kernel."""
        st = self._stats(normalized_acceleration.device)["acceleration"]
        acceleration = normalized_acceleration * st["std"] + st["mean"]
        last = position_sequence[:, -1]
        return last + ((last - position_sequence[:, -2]) + acceleration)

    def _inverse_decoder_postprocessor(self, next_position, position_sequence):
        """Inverse Euler update for the training target (reference :493-517)."""
        st = self._stats(next_position.device)["acceleration"]
        prev = position_sequence[:, -1]
        acceleration = (next_position - prev) - (prev - position_sequence[:, -2])
        return (acceleration - st["mean"]) / st["std"]

    def save(self, path: str = "model.pt"):
        torch.save(self.state_dict(), path)

    def load(self, path: str):
        self.load_state_dict(torch.load(path, map_location=torch.device("cpu")))


def time_diff(position_sequence: torch.Tensor) -> torch.Tensor:
    """Finite difference along the time axis (reference :540-550)."""
    return (position_sequence[:, 1:] - position_sequence[:, :-1]).contiguous()
