"""TEST INFRASTRUCTURE ONLY -- stand-in for the two torch_geometric symbols the reference
imports (`MessagePassing`, `radius_graph`), so that the UNMODIFIED reference under
/root/reference can be imported and run on CPU in the build container to generate the
golden fixtures in tests/golden/ (see oracle/make_golden.py).

torch_geometric / torch_cluster are third-party, absent from /root/reference and not pinned
by it (requirements.txt:1-7; README.md:19-25 suggests the wheels for torch 2.5.0+cu118).
Behaviour restated here (recalled from their published source):

* MessagePassing.propagate (non-fused path): j, i = edge_index[0], edge_index[1];
  `*_i` args = x.index_select(0, i), `*_j` args = x.index_select(0, j); out = message(...);
  aggregate = zeros(N, H).scatter_add_(0, i, out) for aggr='add'; update(aggregate, **kwargs)
  where kwargs are the ones handed to propagate (so update() sees the INPUT edge_features:
  this is what makes the reference's edge residual e' = 2e, graph_network.py:173-176).
* radius_graph: torch_cluster CUDA rule, see oracle/radius_oracle.c.
"""
import inspect
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.dirname(
    os.path.abspath(__file__))))))
from radius import radius_graph_np  # noqa: E402  (oracle/radius.py)


class MessagePassing(torch.nn.Module):
    def __init__(self, aggr="add", flow="source_to_target", node_dim=-2):
        super().__init__()
        self.aggr = aggr
        self.flow = flow
        self.node_dim = node_dim

    def propagate(self, edge_index, size=None, **kwargs):
        assert self.flow == "source_to_target"
        j, i = edge_index[0], edge_index[1]
        msg_params = list(inspect.signature(self.message).parameters)
        upd_params = list(inspect.signature(self.update).parameters)[1:]
        num_nodes = None
        msg_kwargs = {}
        for name in msg_params:
            if name.endswith("_i") or name.endswith("_j"):
                base = kwargs[name[:-2]]
                num_nodes = base.shape[0]
                msg_kwargs[name] = base.index_select(0, i if name.endswith("_i") else j)
            else:
                msg_kwargs[name] = kwargs[name]
        out = self.message(**msg_kwargs)
        if num_nodes is None:
            num_nodes = int(i.max()) + 1 if i.numel() else 0
        if self.aggr in ("add", "sum"):
            agg = out.new_zeros((num_nodes,) + tuple(out.shape[1:]))
            index = i.view(-1, *([1] * (out.dim() - 1))).expand_as(out)
            agg = agg.scatter_add_(0, index, out)
        else:
            raise NotImplementedError(self.aggr)
        upd_kwargs = {name: kwargs[name] for name in upd_params if name in kwargs}
        return self.update(agg, **upd_kwargs)

    def message(self, x_j):
        return x_j

    def update(self, inputs):
        return inputs


def radius_graph(x, r, batch=None, loop=False, max_num_neighbors=32,
                 flow="source_to_target", num_workers=1, batch_size=None):
    assert flow == "source_to_target"
    if x.shape[0] == 0:
        # the reference's test_empty_positions expects a ValueError; real torch_cluster's
        # behaviour for empty input is not pinned.
        raise ValueError("radius_graph: empty input")
    pos = x.detach().cpu().to(torch.float32).numpy()
    if batch is None:
        n_per_example = [pos.shape[0]]
    else:
        b = batch.detach().cpu().numpy()
        assert np.all(np.diff(b) >= 0), "batch vector must be sorted"
        n_per_example = np.bincount(b).tolist()
    ei = radius_graph_np(pos, float(r), n_per_example, loop=loop,
                         max_num_neighbors=max_num_neighbors)
    return torch.from_numpy(ei).to(x.device)
