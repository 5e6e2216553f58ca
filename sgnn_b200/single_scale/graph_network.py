"""Encode-Process-Decode network with the reference's module API and checkpoint keys
(sgnn/single_scale/graph_network.py), executed by the sm_100a kernels of libsgnn_b200.

Class names, constructor signatures, forward signatures and state_dict key names
("NN-i", "node_fn.0.NN-0.weight", "gnn_stacks.{l}.edge_fn.1.weight", ...) match the reference
so its checkpoints load unchanged.  There is no PyTorch/CPU fallback: forward() on a CPU
tensor raises.

The fused per-step path used by LearnedSimulator lives in sgnn_b200/engine.py; the forward()
methods here serve callers that drive the modules directly with explicit feature tensors.
"""
from typing import List

import torch
import torch.nn as nn

from .. import _lib as L
from ..blob import params_version
from ..engine import Csr, EdgeEncoderPlan, EdgeLayerPlan, NodePlan, TILE

_PRECISION = "tf32"


def set_default_precision(precision: str):
    """'tf32' (tcgen05 tensor cores) or 'fp32' (FFMA parity mode) for module-level forwards."""
    global _PRECISION
    if precision not in ("tf32", "fp32"):
        raise ValueError(precision)
    _PRECISION = precision


def build_mlp(input_size: int, hidden_layer_sizes: List[int], output_size: int = None,
              output_activation: nn.Module = nn.Identity,
              activation: nn.Module = nn.ReLU) -> nn.Module:
    """Sequential of Linear "NN-i" / activation "Act-i" pairs (reference build_mlp,
    graph_network.py:7-45); the module names are part of the checkpoint format."""
    sizes = [input_size] + list(hidden_layer_sizes)
    if output_size:
        sizes.append(output_size)
    mlp = nn.Sequential()
    last = len(sizes) - 2
    for i in range(len(sizes) - 1):
        mlp.add_module(f"NN-{i}", nn.Linear(sizes[i], sizes[i + 1]))
        mlp.add_module(f"Act-{i}", (output_activation if i == last else activation)())
    return mlp


class _PlanCache:
    """Rebuilds kernel plans when parameters change (optimizer step, load_state_dict, .to())."""

    def __init__(self):
        self._key, self._plans = None, None

    def get(self, module: nn.Module, build):
        key = (_PRECISION, params_version(module.parameters()))
        if key != self._key:
            self._plans, self._key = build(_PRECISION == "tf32"), key
        return self._plans


def _f32c(t, name):
    L.require_cuda(t, name)
    return t.detach().float().contiguous()


def _rows_forward(plan: NodePlan, x: torch.Tensor, out_w: int, **kw):
    out = torch.empty(x.shape[0], out_w, dtype=torch.float32, device=x.device)
    plan.launch(x.shape[0], rows_in=x, rows_ld=x.shape[1], **kw, **{("dec_out" if plan.p.tail == L.TAIL_DECODE else "x_out"): out})
    return out


class Encoder(nn.Module):
    """Node and edge encoders: MLP + LayerNorm each (reference Encoder, graph_network.py:48-111)."""

    def __init__(self, nnode_in_features: int, nnode_out_features: int, nedge_in_features: int,
                 nedge_out_features: int, nmlp_layers: int, mlp_hidden_dim: int):
        super().__init__()
        hidden = [mlp_hidden_dim for _ in range(nmlp_layers)]
        self.node_fn = nn.Sequential(build_mlp(nnode_in_features, hidden, nnode_out_features),
                                     nn.LayerNorm(nnode_out_features))
        self.edge_fn = nn.Sequential(build_mlp(nedge_in_features, hidden, nedge_out_features),
                                     nn.LayerNorm(nedge_out_features))
        self._cache = _PlanCache()

    def forward(self, x: torch.Tensor, edge_features: torch.Tensor):
        x, ef = _f32c(x, "x"), _f32c(edge_features, "edge_features")
        dev = x.device

        def build(tc):
            return (NodePlan(L.PRO_ROWS, x.shape[1], self.node_fn, tc, dev),
                    EdgeEncoderPlan(self.edge_fn, ef.shape[1], tc, dev))

        node_plan, edge_plan = self._cache.get(self, build)
        xo = _rows_forward(node_plan, x, node_plan.p.mlp.out)
        e = ef.shape[0]
        csr = Csr(1, max(e, 1), dev)      # only its capacity is used: explicit features, no gather
        tiles = torch.empty(csr.max_tiles * TILE * edge_plan.out, dtype=torch.float32, device=dev)
        edge_plan.launch(csr, tiles, feat=ef, n_edges=e)
        eo = torch.empty(e, edge_plan.out, dtype=torch.float32, device=dev)
        import ctypes as C
        L.check(L.lib().sgnn_unpack_tiles_to_rows(L.ptr(tiles), None, e, edge_plan.out, C.c_float(1.0),
                                                  L.ptr(eo), C.c_long(edge_plan.out), L.stream_ptr()))
        return xo, eo


class InteractionNetwork(nn.Module):
    """One message-passing layer (reference InteractionNetwork, graph_network.py:114-222):
    edge MLP on [x_target | x_source | e] -> sum at the target -> node MLP on [aggregate | x]
    -> residual.  As in the reference under PyG's propagate, the returned edge latent is
    `edge_features + edge_features` (update() hands back its input, :173-176,:201-222)."""

    def __init__(self, nnode_in: int, nnode_out: int, nedge_in: int, nedge_out: int,
                 nmlp_layers: int, mlp_hidden_dim: int):
        super().__init__()
        hidden = [mlp_hidden_dim for _ in range(nmlp_layers)]
        self.node_fn = nn.Sequential(build_mlp(nnode_in + nedge_out, hidden, nnode_out),
                                     nn.LayerNorm(nnode_out))
        self.edge_fn = nn.Sequential(build_mlp(nnode_in + nnode_in + nedge_in, hidden, nedge_out),
                                     nn.LayerNorm(nedge_out))
        self._nnode_in, self._nedge_in = nnode_in, nedge_in
        self._cache = _PlanCache()

    def forward(self, x: torch.Tensor, edge_index: torch.Tensor, edge_features: torch.Tensor):
        import ctypes as C
        x, e = _f32c(x, "x"), _f32c(edge_features, "edge_features")
        L.require_cuda(edge_index, "edge_index")
        dev, n = x.device, x.shape[0]
        nn_in, ne_in = self._nnode_in, self._nedge_in

        def build(tc):
            return (NodePlan(L.PRO_ROWS, nn_in, None, tc, dev, nnode_next=nn_in, pq_from=self.edge_fn),
                    EdgeLayerPlan(self.edge_fn, nn_in, ne_in, 1.0, tc, dev),
                    NodePlan(L.PRO_AGG_X, self.edge_fn[1].normalized_shape[0] + nn_in, self.node_fn,
                             tc, dev, residual=True))

        pq_plan, edge_plan, node_plan = self._cache.get(self, build)
        csr, perm = Csr.from_edge_index(edge_index, n)
        n_e = edge_index.shape[1]
        f32 = dict(dtype=torch.float32, device=dev)
        tiles = torch.empty(csr.max_tiles * TILE * ne_in, **f32)
        L.check(L.lib().sgnn_pack_rows_to_tiles(L.ptr(e), C.c_long(ne_in), L.ptr(perm), n_e, ne_in,
                                                int(edge_plan.tc), L.ptr(tiles), L.stream_ptr()))
        pq = torch.empty(n, 2 * edge_plan.hidden, **f32)
        pq_plan.launch(n, rows_in=x, rows_ld=nn_in, pq_out=pq)
        agg = torch.empty(n, edge_plan.out, **f32)
        head = torch.empty(csr.max_tiles, edge_plan.out, **f32)
        edge_plan.launch(csr, tiles, pq, agg, head)
        x_out = torch.empty(n, node_plan.p.mlp.out, **f32)
        node_plan.launch(n, agg=agg, head=head, rowptr=csr.rowptr, agg_w=edge_plan.out, x_in=x,
                         x_w=nn_in, x_out=x_out)
        return x_out, edge_features + edge_features


class Processor(nn.Module):
    """Stack of unshared InteractionNetworks (reference Processor, graph_network.py:225-293)."""

    def __init__(self, nnode_in: int, nnode_out: int, nedge_in: int, nedge_out: int,
                 nmessage_passing_steps: int, nmlp_layers: int, mlp_hidden_dim: int):
        super().__init__()
        self.gnn_stacks = nn.ModuleList([
            InteractionNetwork(nnode_in=nnode_in, nnode_out=nnode_out, nedge_in=nedge_in,
                               nedge_out=nedge_out, nmlp_layers=nmlp_layers,
                               mlp_hidden_dim=mlp_hidden_dim)
            for _ in range(nmessage_passing_steps)])

    def forward(self, x: torch.Tensor, edge_index: torch.Tensor, edge_features: torch.Tensor):
        for gnn in self.gnn_stacks:
            x, edge_features = gnn(x, edge_index, edge_features)
        return x, edge_features


class Decoder(nn.Module):
    """MLP without LayerNorm on the final node latents (reference Decoder, :296-333)."""

    def __init__(self, nnode_in: int, nnode_out: int, nmlp_layers: int, mlp_hidden_dim: int):
        super().__init__()
        self.node_fn = build_mlp(nnode_in, [mlp_hidden_dim for _ in range(nmlp_layers)], nnode_out)
        self._cache = _PlanCache()

    def forward(self, x: torch.Tensor):
        x = _f32c(x, "x")
        plan = self._cache.get(self, lambda tc: NodePlan(L.PRO_ROWS, x.shape[1], None, tc, x.device,
                                                         decoder=self.node_fn))
        return _rows_forward(plan, x, plan.p.dec.out)


class EncodeProcessDecode(nn.Module):
    """Encoder -> Processor -> Decoder (reference EncodeProcessDecode, :336-406)."""

    def __init__(self, nnode_in_features: int, nnode_out_features: int, nedge_in_features: int,
                 latent_dim: int, nmessage_passing_steps: int, nmlp_layers: int, mlp_hidden_dim: int):
        super().__init__()
        self._encoder = Encoder(nnode_in_features=nnode_in_features, nnode_out_features=latent_dim,
                                nedge_in_features=nedge_in_features, nedge_out_features=latent_dim,
                                nmlp_layers=nmlp_layers, mlp_hidden_dim=mlp_hidden_dim)
        self._processor = Processor(nnode_in=latent_dim, nnode_out=latent_dim, nedge_in=latent_dim,
                                    nedge_out=latent_dim, nmessage_passing_steps=nmessage_passing_steps,
                                    nmlp_layers=nmlp_layers, mlp_hidden_dim=mlp_hidden_dim)
        self._decoder = Decoder(nnode_in=latent_dim, nnode_out=nnode_out_features,
                                nmlp_layers=nmlp_layers, mlp_hidden_dim=mlp_hidden_dim)

    def forward(self, x: torch.Tensor, edge_index: torch.Tensor, edge_features: torch.Tensor):
        x, edge_features = self._encoder(x, edge_features)
        x, edge_features = self._processor(x, edge_index, edge_features)
        return self._decoder(x)
