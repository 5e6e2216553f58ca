"""Training path of the Encode-Process-Decode networks (forward WITH autograd + backward).

Inference runs the fused tile kernels (engine.py).  Training needs the intermediate activations
and their gradients, so it is laid out differently:

  * the message-passing glue -- the x_i / x_j gathers, the sum-aggregation and their backward
    passes (the reference's index_select x2 + scatter_add atomics inside PyG's propagate,
    sgnn/single_scale/graph_network.py:173) -- runs on the deterministic CUDA kernels of
    csrc/graph_ops.cu through torch.autograd.Function wrappers (no atomics anywhere: the backward of
    the source gather is a segmented sum over a source-sorted permutation);
  * the [E,3H] and [N,2H] concat buffers of the reference never exist: the first Linear of each
    MLP is applied per column block (P = x W_a^T + b, Q = x W_b^T per NODE, e W_c^T per edge);
  * neighbour search and the node/edge features come from the sm_100a kernels (no gradient flows
    into positions, as in the reference);
  * the dense GEMMs and LayerNorm use cuBLAS / ATen through torch on the very same nn.Linear /
    nn.LayerNorm parameter objects, so autograd produces the parameter gradients and the
    reference's optimizer code (Adam, train.py:199,271-273) runs unchanged.

There is no CPU path: every entry raises on non-CUDA tensors.
"""
import ctypes as C
from typing import List, Optional

import torch
import torch.nn.functional as F

from . import _lib as L
from .blob import linears_of


class TrainGraph:
    """Receiver-sorted edge list (int32) with E known on the host, plus -- built on first use --
    the source-sorted permutation needed by the backward of the source gather."""

    def __init__(self, rowptr: torch.Tensor, src: torch.Tensor, tgt: torch.Tensor, n_nodes: int,
                 n_edges: int):
        self.rowptr, self.n, self.e = rowptr, n_nodes, n_edges
        self.src, self.tgt = src[:n_edges], tgt[:n_edges]
        self._by_src = None

    @staticmethod
    def from_csr(csr, copy: bool = False) -> "TrainGraph":
        """copy=True: own copies of the edge list (the autograd Functions read it again at backward time; a
        persistent neighbour-search buffer would be overwritten by a second forward before that backward)."""
        e = int(csr.rowptr[csr.n_nodes].item())   # the one host sync of a training step
        if copy:
            return TrainGraph(csr.rowptr[:csr.n_nodes + 1].clone(), csr.src[:e].clone(), csr.tgt[:e].clone(),
                              csr.n_nodes, e)
        return TrainGraph(csr.rowptr, csr.src, csr.tgt, csr.n_nodes, e)

    def by_source(self):
        """(perm, rowptr_src): edges stably sorted by source; segment j = edges leaving node j."""
        if self._by_src is None:
            src = self.src.long()
            perm = torch.sort(src, stable=True).indices.int()
            counts = torch.bincount(src, minlength=self.n)
            rp = torch.zeros(self.n + 1, dtype=torch.int32, device=src.device)
            rp[1:] = torch.cumsum(counts, 0).int()
            self._by_src = (perm, rp)
        return self._by_src


def _f32(t):
    L.require_cuda(t, "tensor")
    return t.contiguous().float()


def segment_sum_rows(m: torch.Tensor, rowptr: torch.Tensor, n_rows: int, perm=None) -> torch.Tensor:
    m = _f32(m)
    out = torch.empty(n_rows, m.shape[1], dtype=torch.float32, device=m.device)
    L.check(L.lib().sgnn_segment_sum_rows(L.ptr(m), L.ptr(rowptr), L.ptr(perm), n_rows, m.shape[1],
                                          L.ptr(out), L.stream_ptr()))
    return out


def gather_rows(rows: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    rows = _f32(rows)
    out = torch.empty(idx.numel(), rows.shape[1], dtype=torch.float32, device=rows.device)
    L.check(L.lib().sgnn_gather_rows(L.ptr(rows), C.c_long(rows.shape[1]), L.ptr(idx), idx.numel(),
                                     rows.shape[1], L.ptr(out), L.stream_ptr()))
    return out


class _GatherAdd(torch.autograd.Function):
    """out[e] = P[tgt[e]] + Q[src[e]]."""

    @staticmethod
    def forward(ctx, p, q, graph: TrainGraph):
        p, q = _f32(p), _f32(q)
        ctx.graph = graph
        out = torch.empty(graph.e, p.shape[1], dtype=torch.float32, device=p.device)
        L.check(L.lib().sgnn_gather_add_rows(L.ptr(p), C.c_long(p.shape[1]), L.ptr(q), C.c_long(q.shape[1]),
                                             L.ptr(graph.tgt), L.ptr(graph.src), graph.e, p.shape[1],
                                             L.ptr(out), L.stream_ptr()))
        return out

    @staticmethod
    def backward(ctx, dout):
        g = ctx.graph
        dout = _f32(dout)
        dp = segment_sum_rows(dout, g.rowptr, g.n)                     # edges are receiver-sorted
        perm, rp = g.by_source()
        dq = segment_sum_rows(dout, rp, g.n, perm)
        return dp, dq, None


class _SegmentSum(torch.autograd.Function):
    """agg[i] = sum of m over the CSR row of receiver i (rows summed in edge order)."""

    @staticmethod
    def forward(ctx, m, graph: TrainGraph):
        ctx.graph = graph
        return segment_sum_rows(m, graph.rowptr, graph.n)

    @staticmethod
    def backward(ctx, dagg):
        return gather_rows(dagg, ctx.graph.tgt), None


class _LayerNorm(torch.autograd.Function):
    """nn.LayerNorm over rows of 32 / 64 / 128 floats on csrc/train_norm.cu (forward keeps mean / rstd,
    backward is one streaming pass + a fixed-order column reduction: deterministic dgamma / dbeta)."""

    @staticmethod
    def forward(ctx, x, gamma, beta, eps: float):
        x, gamma, beta = _f32(x), _f32(gamma), _f32(beta)
        n, w = x.shape
        y = torch.empty_like(x)
        mean = torch.empty(n, dtype=torch.float32, device=x.device)
        rstd = torch.empty(n, dtype=torch.float32, device=x.device)
        L.check(L.lib().sgnn_layer_norm_fwd(L.ptr(x), C.c_long(n), w, L.ptr(gamma), L.ptr(beta), C.c_float(eps),
                                            L.ptr(y), L.ptr(mean), L.ptr(rstd), L.stream_ptr()))
        ctx.save_for_backward(x, gamma, mean, rstd)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, gamma, mean, rstd = ctx.saved_tensors
        dy = _f32(dy)
        n, w = x.shape
        dx = torch.empty_like(x)
        dgamma, dbeta = torch.empty_like(gamma), torch.empty_like(gamma)
        nbytes = L.lib().sgnn_layer_norm_bwd_workspace_bytes(w)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=x.device)
        L.check(L.lib().sgnn_layer_norm_bwd(L.ptr(dy), L.ptr(x), L.ptr(mean), L.ptr(rstd), L.ptr(gamma), C.c_long(n),
                                            w, L.ptr(dx), L.ptr(dgamma), L.ptr(dbeta), L.ptr(ws),
                                            C.c_size_t(nbytes), L.stream_ptr()), launches=2)
        return dx, dgamma, dbeta, None


_wgrad_ws = {}          # (device, stream) -> persistent workspace: kernels of one stream run in order, so reuse is safe


def _wgrad_workspace(nbytes: int, device) -> torch.Tensor:
    key = (device, torch.cuda.current_stream(device).cuda_stream)
    ws = _wgrad_ws.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = torch.empty(max(nbytes, 1 << 24), dtype=torch.uint8, device=device)
        _wgrad_ws[key] = ws
    return ws


class _Linear(torch.autograd.Function):
    """y = x W^T + b.  Forward and input gradient are plain library GEMMs; the weight and bias gradients
    -- 64 x 64 outputs summed over up to 200 k rows, which cuBLAS runs inside one or two CTAs -- come from
    csrc/train_linear.cu (rows split over the grid, fixed-order reduction: deterministic)."""

    @staticmethod
    def forward(ctx, x, w, b):
        ctx.save_for_backward(x, w)
        ctx.has_bias = b is not None
        return F.linear(x, w, b)

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        if dy.dtype != torch.float32 or not dy.is_contiguous():
            dy = _f32(dy)
        dx = dy @ w if ctx.needs_input_grad[0] else None
        dw = db = None
        if ctx.needs_input_grad[1] or (ctx.has_bias and ctx.needs_input_grad[2]):
            xc = x if (x.dtype == torch.float32 and x.is_contiguous()) else _f32(x)
            dw, db = _wgrad(dy, xc, ctx.has_bias)
        return dx, dw, db


_wgrad_fn = None


def _wgrad_fns():
    """The two C entries with their argtypes declared once (this call sits on the backward path of 41 Linears
    per step: no per-call ctypes object construction)."""
    global _wgrad_fn
    if _wgrad_fn is None:
        lib = L.lib()
        f, g = lib.sgnn_linear_wgrad, lib.sgnn_linear_wgrad_workspace_bytes
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_long, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p,
                      C.c_size_t, C.c_void_p]
        g.argtypes = [C.c_long, C.c_int, C.c_int]
        _wgrad_fn = (f, g)
    return _wgrad_fn


def linear(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor] = None) -> torch.Tensor:
    """F.linear on [rows, in] activations with the weight gradient on the CUDA path above."""
    L.require_cuda(x, "linear input")       # no CPU path: the training step runs on the GPU or not at all
    if x.dim() != 2 or not (torch.is_grad_enabled() and (w.requires_grad or (b is not None and b.requires_grad))):
        return F.linear(x, w, b)
    return _Linear.apply(x, w, b)


def _wgrad(dy: torch.Tensor, x: torch.Tensor, bias: bool):
    """(dW [out, in], db [out] or None) of y = x W^T (+ b) for contiguous fp32 dy [rows, out], x [rows, in]."""
    n, k = x.shape
    m = dy.shape[1]
    dev = dy.device
    dw = torch.empty(m, k, dtype=torch.float32, device=dev)
    db = torch.empty(m, dtype=torch.float32, device=dev) if bias else None
    fn, fn_bytes = _wgrad_fns()
    ws = _wgrad_workspace(fn_bytes(n, m, k), dev)
    L.check(fn(dy.data_ptr(), x.data_ptr(), n, m, k, dw.data_ptr(), db.data_ptr() if bias else None,
               ws.data_ptr(), ws.numel(), torch.cuda.current_stream().cuda_stream), launches=2)
    return dw, db


class _EdgePre(torch.autograd.Function):
    """Pre-activation of the edge MLP's first Linear on cat([x_i, x_j, e]) without the concatenation
    (graph_network.py:173,197): pre[k] = e[k] Wc^T + b + (x Wa^T)[tgt[k]] + (x Wb^T)[src[k]] with
    W = [Wa | Wb | Wc].  One node-level GEMM for both x terms, the gather added in place; the backward
    sums dpre per receiver and per sender into the two halves of one [N, 2H] matrix, so x sees one
    input-gradient GEMM and W one assembled gradient instead of three sliced ones."""

    @staticmethod
    def forward(ctx, x, e, w, b, graph: TrainGraph):
        x, e = _f32(x), _f32(e)
        ni, h = x.shape[1], w.shape[0]
        wst = torch.cat([w[:, :ni], w[:, ni:2 * ni]], 0)              # [2H, ni]: rows of Wa, then rows of Wb
        pq = x @ wst.t()                                              # [N, 2H] = [P | Q]
        wc = w[:, 2 * ni:]
        pre = torch.addmm(b, e, wc.t()) if b is not None else e @ wc.t()
        L.check(L.lib().sgnn_gather_add_rows_acc(L.ptr(pq), C.c_long(2 * h), C.c_void_p(pq.data_ptr() + 4 * h),
                                                 C.c_long(2 * h), L.ptr(graph.tgt), L.ptr(graph.src), graph.e, h,
                                                 L.ptr(pre), L.stream_ptr()))
        ctx.save_for_backward(x, e, w, wst)
        ctx.graph, ctx.has_bias = graph, b is not None
        return pre

    @staticmethod
    def backward(ctx, dpre):
        x, e, w, wst = ctx.saved_tensors
        g = ctx.graph
        if dpre.dtype != torch.float32 or not dpre.is_contiguous():
            dpre = _f32(dpre)
        ni, h = x.shape[1], w.shape[0]
        dwc, db = _wgrad(dpre, e, ctx.has_bias)
        de = dpre @ w[:, 2 * ni:] if ctx.needs_input_grad[1] else None
        dpq = torch.empty(g.n, 2 * h, dtype=torch.float32, device=dpre.device)
        perm, rp = g.by_source()
        lib, st = L.lib(), L.stream_ptr()
        L.check(lib.sgnn_segment_sum_rows_ld(L.ptr(dpre), L.ptr(g.rowptr), None, g.n, h, L.ptr(dpq),
                                             C.c_long(2 * h), st))
        L.check(lib.sgnn_segment_sum_rows_ld(L.ptr(dpre), L.ptr(rp), L.ptr(perm), g.n, h,
                                             C.c_void_p(dpq.data_ptr() + 4 * h), C.c_long(2 * h), st))
        dwst, _ = _wgrad(dpq, x, False)                               # [2H, ni]
        dx = dpq @ wst if ctx.needs_input_grad[0] else None
        dw = torch.cat([dwst[:h], dwst[h:], dwc], 1)
        return dx, de, dw, db, None


def edge_pre(x, e, w, b, graph: TrainGraph):
    return _EdgePre.apply(x, e, w, b, graph)


def apply_mlp(seq_mlp, h, skip_first_linear: bool = False):
    """A build_mlp Sequential (Linear / activation children) with its Linears through linear()."""
    skip = skip_first_linear
    for m in seq_mlp.children():
        if isinstance(m, torch.nn.Linear):
            if skip:
                skip = False
                continue
            h = linear(h, m.weight, m.bias)
        else:
            h = m(h)
    return h


def layer_norm(ln: torch.nn.LayerNorm, h: torch.Tensor) -> torch.Tensor:
    """ln(h) for a reference nn.LayerNorm module on [rows, width] activations."""
    w = h.shape[-1]
    if h.dim() != 2 or not ln.elementwise_affine or ln.bias is None or tuple(ln.normalized_shape) != (w,) \
            or not L.lib().sgnn_layer_norm_supported(w):
        return ln(h)            # widths outside {32, 64, 128}: ATen (still CUDA; training glue only)
    return _LayerNorm.apply(h, ln.weight, ln.bias, float(ln.eps))


def mlp_ln(seq: torch.nn.Sequential, h: torch.Tensor) -> torch.Tensor:
    """Sequential(build_mlp, LayerNorm) as the encoders use it."""
    return layer_norm(seq[1], apply_mlp(seq[0], h))


def gather_add(p, q, graph):
    return _GatherAdd.apply(p, q, graph)


def segment_sum(m, graph):
    return _SegmentSum.apply(m, graph)


# ---------------------------------------------------------------------------------------------
def node_features(seq: torch.Tensor, vel_mean, vel_std, wall_max: float, wall_div: float,
                  extra: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[N, (S-1)*dim + 1 (+ extra columns)] node features from the position window."""
    n, s, dim = seq.shape
    k = (s - 1) * dim + 1
    width = k + (0 if extra is None else extra.shape[1])
    out = torch.empty(n, width, dtype=torch.float32, device=seq.device)
    L.check(L.lib().sgnn_node_features(L.ptr(seq), C.c_long(seq.stride(0)), n, s, dim, L.ptr(vel_mean),
                                       L.ptr(vel_std), C.c_float(2.0), C.c_float(wall_max),
                                       C.c_float(wall_div), L.ptr(out), C.c_long(width), L.stream_ptr()))
    if extra is not None:
        out = torch.cat([out[:, :k], extra], dim=1)     # keeps the embedding's autograd edge
    return out


def edge_features(last_pos: torch.Tensor, pos_stride: int, dim: int, radius: float,
                  graph: TrainGraph) -> torch.Tensor:
    out = torch.empty(graph.e, dim + 1, dtype=torch.float32, device=graph.src.device)
    L.check(L.lib().sgnn_edge_features(C.c_void_p(last_pos.data_ptr()), C.c_long(pos_stride), dim,
                                       C.c_float(radius), L.ptr(graph.src), L.ptr(graph.tgt), graph.e,
                                       L.ptr(out), L.stream_ptr()))
    return out


def _rest_of_mlp(seq_mlp, h):
    """Apply everything after the first Linear of a build_mlp Sequential (Act-0, NN-1, ...)."""
    return apply_mlp(seq_mlp, h, skip_first_linear=True)


def interaction_block(edge_fn, node_fn, x, e, graph: TrainGraph):
    """One InteractionNetwork / G2M / M2M / M2G block with autograd (graph_network.py:150-222).
    Returns (x + node update, e + e): the reference's update() hands back its input edge latent."""
    nn_in = x.shape[1]
    w1 = linears_of(edge_fn[0])[0]
    if w1.weight.shape[1] > 2 * nn_in:
        pre = edge_pre(x, e, w1.weight, w1.bias, graph)
    else:
        p = linear(x, w1.weight[:, :nn_in], w1.bias)                # x_target block (+ bias)
        q = linear(x, w1.weight[:, nn_in:2 * nn_in])                # x_source block
        pre = linear(e, w1.weight[:, 2 * nn_in:]) + gather_add(p, q, graph)
    msg = layer_norm(edge_fn[1], _rest_of_mlp(edge_fn[0], pre))   # ... -> LayerNorm
    agg = segment_sum(msg, graph)
    v1 = linears_of(node_fn[0])[0]
    pre_n = linear(torch.cat([agg, x], 1), v1.weight, v1.bias)    # the reference's cat([aggregated, x]) + Linear
    upd = layer_norm(node_fn[1], _rest_of_mlp(node_fn[0], pre_n))
    return x + upd, e + e


def epd_forward(node_encoder, edge_encoders: List, blocks: List, decoder, nf, efs: List,
                graphs: List[TrainGraph]):
    """blocks: list of (edge_fn, node_fn, edge_set_index).  Returns the decoder output [N, dim+1]."""
    x = mlp_ln(node_encoder, nf)
    es = [mlp_ln(enc, ef) for enc, ef in zip(edge_encoders, efs)]
    for edge_fn, node_fn, k in blocks:
        x, es[k] = interaction_block(edge_fn, node_fn, x, es[k], graphs[k])
    return apply_mlp(decoder, x) if isinstance(decoder, torch.nn.Sequential) else decoder(x)
