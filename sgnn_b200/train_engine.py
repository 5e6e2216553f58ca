"""Training step of the single-scale simulator as hand-written sm_100a kernels under ONE CUDA graph
(reference: sgnn/single_scale/train.py:230-280 -- noise, predict_accelerations, loss, backward, Adam --
and the autograd of sgnn/single_scale/graph_network.py:21-58,86-96,139-222).

train_ops.py runs the same step through torch.autograd with library GEMMs and one host read per step
(the edge count), ~380 launches issued from Python.  Here the backward pass is written out by hand:

  * every Linear / Linear-ReLU-Linear-LayerNorm block is one launch of csrc/train_mlp.cu (tcgen05 on
    hi/lo-split TF32 operands = fp32-grade products, or plain TF32), forward keeping h1, the LayerNorm
    input and its statistics, backward chaining dy W2 -> ReLU mask -> dpre W1;
  * LayerNorm backward (+ the gather of the aggregation's backward), weight / bias gradients and the
    segmented sums are the deterministic kernels of train_norm.cu / train_linear.cu / graph_ops.cu, all
    taking the edge count from device memory (rowptr[N]);
  * gradients land in ONE flat buffer (every p.grad is a view of it): the data-parallel allreduce is a
    single NCCL call inside the graph, Adam is torch's fused, capturable kernel on the same tensors;
  * buffers are sized by the capacity N * 20 edges, so a step never allocates and never reads the host:
    pack weights -> neighbour search -> forward -> loss -> backward -> allreduce -> Adam replays as one graph.

Supported: the single-scale network with nmlp_layers = 1, one particle type, latent = hidden = 64
(what train._get_simulator builds for BASELINE configs[2]); anything else keeps the autograd path.
"""
import ctypes as C
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from . import _lib as L
from . import noise_utils
from .blob import linears_of, pad32
from .engine import RadiusGraph
from .single_scale.learned_simulator import MAX_NUM_NEIGHBORS, example_ptr_of

TPRO_ROWS, TPRO_ROWS2, TPRO_EDGE_FEATURES = 0, 1, 2


class TrainFwd(C.Structure):
    _fields_ = [("max_rows", C.c_int), ("n_rows_dev", C.c_void_p), ("prologue", C.c_int),
                ("a0", C.c_void_p), ("ld0", C.c_long), ("k0", C.c_int), ("scale0", C.c_float),
                ("a1", C.c_void_p), ("ld1", C.c_long), ("k1", C.c_int),
                ("pos", C.c_void_p), ("pos_stride", C.c_long), ("dim", C.c_int), ("radius", C.c_float),
                ("src", C.c_void_p), ("tgt", C.c_void_p), ("feat_out", C.c_void_p),
                ("add_p", C.c_void_p), ("add_q", C.c_void_p), ("ld_add", C.c_long),
                ("kpad", C.c_int), ("hidden", C.c_int), ("out", C.c_int), ("n_linear", C.c_int),
                ("has_ln", C.c_int), ("split", C.c_int),
                ("w1_hi", C.c_void_p), ("w1_lo", C.c_void_p), ("w2_hi", C.c_void_p), ("w2_lo", C.c_void_p),
                ("b1", C.c_void_p), ("b1_cols", C.c_int), ("b2", C.c_void_p), ("gamma", C.c_void_p),
                ("beta", C.c_void_p), ("res", C.c_void_p), ("ld_res", C.c_long),
                ("h1", C.c_void_p), ("y2", C.c_void_p), ("mean", C.c_void_p), ("rstd", C.c_void_p),
                ("y", C.c_void_p), ("ld_y", C.c_long)]


class TrainBwd(C.Structure):
    _fields_ = [("max_rows", C.c_int), ("n_rows_dev", C.c_void_p), ("dy", C.c_void_p), ("ld_dy", C.c_long),
                ("k_dy", C.c_int), ("n_linear", C.c_int), ("hidden", C.c_int), ("h1", C.c_void_p),
                ("dpre", C.c_void_p), ("w2t_hi", C.c_void_p), ("w2t_lo", C.c_void_p), ("want_dx", C.c_int),
                ("kx", C.c_int), ("w1t_hi", C.c_void_p), ("w1t_lo", C.c_void_p),
                ("dx0", C.c_void_p), ("ld_dx0", C.c_long), ("k0", C.c_int), ("acc0", C.c_int),
                ("dx1", C.c_void_p), ("ld_dx1", C.c_long), ("k1", C.c_int), ("acc1", C.c_int),
                ("add1", C.c_void_p), ("ld_add1", C.c_long), ("split", C.c_int)]


class PackDesc(C.Structure):
    _fields_ = [("src", C.c_void_p), ("stride_n", C.c_long), ("stride_k", C.c_long), ("n", C.c_int), ("k", C.c_int),
                ("row0", C.c_int), ("col0", C.c_int), ("n_pad", C.c_int), ("pad_", C.c_int), ("scale", C.c_float),
                ("pad2_", C.c_int), ("dst_off", C.c_long)]


class WgradItem(C.Structure):
    _fields_ = [("dy", C.c_void_p), ("x", C.c_void_p), ("dw", C.c_void_p), ("db", C.c_void_p), ("ld_dy", C.c_long),
                ("ld_x", C.c_long), ("ld_dw", C.c_long), ("scale", C.c_float), ("pad_", C.c_int)]


class WgradBatch(C.Structure):
    _fields_ = [("item", WgradItem * 6), ("n_items", C.c_int), ("split", C.c_int), ("max_rows", C.c_long),
                ("n_rows_dev", C.c_void_p), ("workspace", C.c_void_p), ("workspace_bytes", C.c_size_t)]


def supports(sim) -> bool:
    """True when `sim` is a single-scale LearnedSimulator this engine covers (module docstring)."""
    epd = getattr(sim, "_encode_process_decode", None)
    if epd is None or getattr(sim, "_nparticle_types", 2) != 1:
        return False
    h = sim._latent_dim
    if h != 64:                       # the P | Q product is one N = 2h accumulator (<= 128 columns per launch)
        return False
    mlps = [epd._encoder.node_fn[0], epd._encoder.edge_fn[0], epd._decoder.node_fn]
    for g in epd._processor.gnn_stacks:
        mlps += [g.edge_fn[0], g.node_fn[0]]
    for m in mlps:
        lins = linears_of(m)
        if len(lins) != 2 or lins[0].out_features != h or any(l.bias is None for l in lins):
            return False
    return True


class _Images:
    """Plans the weight images of a step: add() returns the float offset of an image; blocks() fill it."""

    def __init__(self):
        self.descs: List[PackDesc] = []
        self.floats = 0

    def image(self, n_pad: int, kpad: int) -> int:
        off = self.floats
        self.floats += (n_pad * kpad + 255) // 256 * 256          # 1024-byte aligned images
        return off

    def block(self, off: int, w: torch.Tensor, n: int, k: int, stride_n: int, stride_k: int, n_pad: int,
              elem0: int = 0, row0: int = 0, col0: int = 0, scale: float = 1.0):
        d = PackDesc()
        d.src = w.data_ptr() + 4 * elem0
        d.stride_n, d.stride_k, d.n, d.k, d.row0, d.col0, d.n_pad = stride_n, stride_k, n, k, row0, col0, n_pad
        d.scale, d.dst_off = scale, off
        self.descs.append(d)

    def plain(self, w: torch.Tensor, rows: slice = None, cols: slice = None, scale: float = 1.0, n_pad=None) -> int:
        """image of W[rows, cols] (N = rows, K = cols) of a contiguous [out, in] weight"""
        ld = w.shape[1]
        r0, r1 = (0, w.shape[0]) if rows is None else (rows.start, rows.stop)
        c0, c1 = (0, ld) if cols is None else (cols.start, cols.stop)
        n, k = r1 - r0, c1 - c0
        n_pad = (n + 15) // 16 * 16 if n_pad is None else n_pad
        off = self.image(n_pad, pad32(k))
        self.block(off, w, n, k, ld, 1, n_pad, elem0=r0 * ld + c0, scale=scale)
        return off

    def transposed(self, w: torch.Tensor, rows: slice = None, cols: slice = None, scale: float = 1.0) -> int:
        """image of W[rows, cols]^T (N = cols, K = rows)"""
        ld = w.shape[1]
        r0, r1 = (0, w.shape[0]) if rows is None else (rows.start, rows.stop)
        c0, c1 = (0, ld) if cols is None else (cols.start, cols.stop)
        n, k = c1 - c0, r1 - r0
        n_pad = (n + 15) // 16 * 16
        off = self.image(n_pad, pad32(k))
        self.block(off, w, n, k, 1, ld, n_pad, elem0=r0 * ld + c0, scale=scale)
        return off


class _Timed:
    def __init__(self, store, name):
        self.store, self.name = store, name

    def __enter__(self):
        if self.store is not None:
            self.ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
            self.ev[0].record()

    def __exit__(self, *a):
        if self.store is not None:
            self.ev[1].record()
            self.store.setdefault(self.name, []).append(self.ev)


class TrainEngine:
    """One training step of `sim` on n particles per rank (module docstring).

        eng = TrainEngine(sim, optimizer, n, device)
        loss = eng.step(batch, noise_std=..., lr=...)        # device scalar, no host sync

    `optimizer` must be torch.optim.Adam(fused=True, capturable=True) over sim.parameters() with a tensor
    learning rate (training.make_optimizer builds it)."""

    def __init__(self, sim, optimizer, n: int, device, precision: str = "fp32", w_pos: float = 1.0,
                 w_strain: float = 1.0, group=None, use_graph: bool = True, distributed: bool = True):
        if not supports(sim):
            raise L.SgnnError("TrainEngine: unsupported model (needs single-scale, nmlp_layers=1, one particle type, "
                              "latent 64)")
        if precision not in ("fp32", "tf32"):
            raise ValueError("precision must be 'fp32' (hi/lo-split TF32 products) or 'tf32'")
        self.sim, self.opt, self.n, self.dev = sim, optimizer, int(n), torch.device(device)
        self.split = 1 if precision == "fp32" else 0
        self.precision = precision
        self.w_pos, self.w_strain = float(w_pos), float(w_strain)
        self.group = group
        # distributed=False: a rank-local engine inside an initialised process group (no gradient allreduce)
        self.world = dist.get_world_size(group) if (distributed and dist.is_initialized()) else 1
        self.use_graph = use_graph
        self.graph = None
        self.dim = sim._particle_dimensions
        self.h = sim._latent_dim
        self.radius = float(sim._connectivity_radius)
        self.lib = L.lib()
        dev, n, h = self.dev, self.n, self.h
        epd = sim._encode_process_decode
        self.layers = list(epd._processor.gnn_stacks)
        nl = len(self.layers)

        # ---- flat gradient buffer: every p.grad is a view of it ----
        self.params = [p for p in sim.parameters() if p.requires_grad]
        self.flat_grad = torch.zeros(sum(p.numel() for p in self.params), dtype=torch.float32, device=dev)
        self._attach_grads()

        # ---- graph / edge buffers at capacity ----
        self.rg = RadiusGraph(n, MAX_NUM_NEIGHBORS, True, dev)
        csr = self.rg.csr
        self.cap = cap = csr.max_edges
        self.n_edges_dev = csr.rowptr.data_ptr() + 4 * n                  # int32 rowptr[n], read by every edge kernel
        self.ws_src = torch.empty(self.lib.sgnn_csr_by_source_workspace_bytes(n), dtype=torch.uint8, device=dev)
        self.perm_src = torch.zeros(cap, dtype=torch.int32, device=dev)
        self.rp_src = torch.zeros(n + 1, dtype=torch.int32, device=dev)

        f = lambda *shape: torch.zeros(*shape, dtype=torch.float32, device=dev)     # noqa: E731
        s = self.s = None
        self.k_node = sim._nnode_in
        # static inputs of a step
        self.window = None                         # [n, S, dim], allocated at the first step (S from the batch)
        self.next_pos = f(n, self.dim)
        self.next_strain = f(n)
        self.noise = None
        self.example_ptr = None
        self.inv_n = f(1)
        self.loss = f(1)
        # activations
        self.k_node_pad = pad32(self.k_node)
        self.nf = f(n, self.k_node_pad)             # zero-padded rows: 16-byte aligned, whole-row requests
        self.ef = f(cap, self.dim + 1)
        mk_n = lambda: {"h1": f(n, h), "y2": f(n, h), "mean": f(n), "rstd": f(n)}           # noqa: E731
        mk_e = lambda: {"h1": f(cap, h), "y2": f(cap, h), "mean": f(cap), "rstd": f(cap)}   # noqa: E731
        self.act_ne, self.act_ee = mk_n(), mk_e()
        self.act_n = [mk_n() for _ in range(nl)]
        self.act_e = [mk_e() for _ in range(nl)]
        self.x = [f(n, h) for _ in range(nl + 1)]
        self.e0 = f(cap, h)
        self.pq = [f(n, 2 * h) for _ in range(nl)]
        self.agg = [f(n, h) for _ in range(nl)]
        self.msg = f(cap, h)
        self.h1_d = f(n, h)
        self.pred = f(n, 4)                        # [:, :dim + 1]; leading dimension 4 keeps rows 16-byte aligned
        assert self.dim + 1 <= 4
        # gradients of activations
        self.dpred = f(n, 4)
        self.dx = [f(n, h), f(n, h)]
        self.dagg = f(n, h)
        # One buffer per use: the weight-gradient kernels run on a side stream and may still read a layer's dy2 / dpre
        # while the main stream is already producing the next layer's.
        self.dy2_n_all = [f(n, h) for _ in range(nl + 1)]           # node blocks, node encoder
        self.dpre_n_all = [f(n, h) for _ in range(nl + 2)]          # + decoder
        self.dy2_e_all = [f(cap, h) for _ in range(nl + 1)]         # edge blocks, edge encoder
        self.dpre_e_all = [f(cap, h) for _ in range(nl + 1)]
        self.dpq_all = [f(n, 2 * h) for _ in range(nl)]
        self.de0 = f(cap, h)
        nb = max(self.lib.sgnn_layer_norm_bwd_workspace_bytes(h),
                 self.lib.sgnn_linear_wgrad_workspace_bytes(C.c_long(cap), 2 * h, 2 * h),
                 self.lib.sgnn_linear_wgrad_workspace_bytes(C.c_long(n), 2 * h, 2 * h))
        self.ws = torch.empty(nb, dtype=torch.uint8, device=dev)          # LayerNorm backward (main stream)
        nb_w = max([nb] + [self.lib.sgnn_linear_wgrad_tc_workspace_bytes(i, r) for i in range(1, 7) for r in (n, cap)])
        self.ws_w = torch.empty(nb_w, dtype=torch.uint8, device=dev)      # weight gradients (side stream)
        self.side = torch.cuda.Stream(device=dev)
        self.overlap_wgrad = True      # weight / bias gradients feed nothing but the optimizer: off the critical path
        self.tc_wgrad = precision == "tf32"
        self._wq = []
        self._declare()
        self._plan_images()

    # ------------------------------------------------------------------ set-up
    def _attach_grads(self):
        off = 0
        for p in self.params:
            k = p.numel()
            view = self.flat_grad[off:off + k].view_as(p)
            if p.grad is None or p.grad.data_ptr() != view.data_ptr():
                p.grad = view
            off += k

    def _declare(self):
        lib = self.lib
        vp, lg, it, fl, sz = C.c_void_p, C.c_long, C.c_int, C.c_float, C.c_size_t
        lib.sgnn_train_pack.argtypes = [vp, it, vp, vp, vp]
        lib.sgnn_csr_by_source.argtypes = [vp, vp, it, it, vp, vp, vp, sz, vp]
        lib.sgnn_train_mlp_fwd.argtypes = [vp, vp]
        lib.sgnn_train_mlp_bwd.argtypes = [vp, vp]
        lib.sgnn_layer_norm_bwd_rows.argtypes = [vp, vp, vp, vp, vp, vp, lg, vp, it, vp, vp, vp, vp, sz, vp]
        lib.sgnn_linear_wgrad_ex.argtypes = [vp, lg, vp, lg, lg, vp, it, it, vp, lg, fl, vp, vp, sz, vp]
        lib.sgnn_segment_sum_rows_ld.argtypes = [vp, vp, vp, it, it, vp, lg, vp]
        lib.sgnn_node_features.argtypes = [vp, lg, it, it, it, vp, vp, fl, fl, fl, vp, lg, vp]

    def _plan_images(self):
        """Offsets of every forward / backward weight image and the descriptor table of the pack kernel."""
        im = _Images()
        h = self.h
        epd = self.sim._encode_process_decode
        self.mlp = {}

        def two(seq_mlp):
            a, b = linears_of(seq_mlp)
            return a, b

        a, b = two(epd._encoder.node_fn[0])
        self.mlp["ne"] = dict(a=a, b=b, ln=epd._encoder.node_fn[1], w1=im.plain(a.weight), w2=im.plain(b.weight),
                              w2t=im.transposed(b.weight))
        a, b = two(epd._encoder.edge_fn[0])
        self.mlp["ee"] = dict(a=a, b=b, ln=epd._encoder.edge_fn[1], w1=im.plain(a.weight), w2=im.plain(b.weight),
                              w2t=im.transposed(b.weight))
        self.mlp_e, self.mlp_n = [], []
        for l, g in enumerate(self.layers):
            a, b = two(g.edge_fn[0])
            w = a.weight                                                  # [h, 3h] = [Wa | Wb | Wc]
            sc = float(2 ** l)                                            # the reference's e' = e + e (graph_network.py:176)
            pq = im.image(2 * h, h)                                       # rows 0..h-1 = Wa, h..2h-1 = Wb
            im.block(pq, w, h, h, 3 * h, 1, 2 * h, elem0=0, row0=0)
            im.block(pq, w, h, h, 3 * h, 1, 2 * h, elem0=h, row0=h)
            pqt = im.image(h, 2 * h)                                      # [Wa; Wb]^T: N = h, K = 2h
            im.block(pqt, w, h, h, 1, 3 * h, h, elem0=0, col0=0)
            im.block(pqt, w, h, h, 1, 3 * h, h, elem0=h, col0=h)
            self.mlp_e.append(dict(a=a, b=b, ln=g.edge_fn[1], scale=sc, pq=pq, pqt=pqt,
                                   w1=im.plain(w, cols=slice(2 * h, 3 * h), scale=sc), w2=im.plain(b.weight),
                                   w2t=im.transposed(b.weight),
                                   w1t=im.transposed(w, cols=slice(2 * h, 3 * h), scale=sc)))
            a, b = two(g.node_fn[0])
            self.mlp_n.append(dict(a=a, b=b, ln=g.node_fn[1], w1=im.plain(a.weight), w2=im.plain(b.weight),
                                   w2t=im.transposed(b.weight), w1t=im.transposed(a.weight)))
        a, b = two(epd._decoder.node_fn)
        self.mlp["dec"] = dict(a=a, b=b, w1=im.plain(a.weight), w2=im.plain(b.weight), w2t=im.transposed(b.weight),
                               w1t=im.transposed(a.weight))
        self.img_hi = torch.zeros(im.floats, dtype=torch.float32, device=self.dev)
        self.img_lo = torch.zeros(im.floats, dtype=torch.float32, device=self.dev)
        arr = (PackDesc * len(im.descs))(*im.descs)
        raw = bytes(arr)
        self.descs = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(self.dev)
        self.n_descs = len(im.descs)
        self._param_ptrs = tuple(p.data_ptr() for p in self.params)

    def _img(self, off):
        return self.img_hi.data_ptr() + 4 * off, self.img_lo.data_ptr() + 4 * off

    # ------------------------------------------------------------------ launches
    profile = None          # set to {} to collect CUDA-event pairs per kernel family (un-captured steps only)

    def _timed(self, name):
        return _Timed(self.profile, name)

    def profile_summary(self):
        """{family: (launch groups, total ms)} of the last profiled step; call after torch.cuda.synchronize()."""
        return {k: (len(v), sum(a.elapsed_time(b) for a, b in v)) for k, v in (self.profile or {}).items()}

    def _fwd(self, tag="fwd", **kw):
        with self._timed(tag):
            self._fwd_launch(**kw)

    def _bwd(self, tag="bwd", **kw):
        with self._timed(tag):
            self._bwd_launch(**kw)

    def _fwd_launch(self, **kw):
        p = TrainFwd()
        p.scale0 = 1.0
        p.split = self.split
        for k, v in kw.items():
            if isinstance(v, torch.Tensor):
                v = v.data_ptr()
            setattr(p, k, v)
        L.check(self.lib.sgnn_train_mlp_fwd(C.addressof(p), self._stream()))

    def _bwd_launch(self, **kw):
        p = TrainBwd()
        p.split = self.split
        for k, v in kw.items():
            if isinstance(v, torch.Tensor):
                v = v.data_ptr()
            setattr(p, k, v)
        L.check(self.lib.sgnn_train_mlp_bwd(C.addressof(p), self._stream()))

    def _stream(self):
        return torch.cuda.current_stream().cuda_stream

    def _wgrad(self, dy, ld_dy, x, ld_x, rows, rows_dev, m, k, dw: torch.Tensor, col0: int, scale: float = 1.0,
               db: Optional[torch.Tensor] = None):
        """Queues dW[:, col0:col0+k] = scale * dy^T x (and db = column sums of dy) into the parameter's .grad view;
        _wgrad_flush() launches the queue."""
        ptr = lambda t: t if isinstance(t, int) else t.data_ptr()      # noqa: E731
        self._wq.append((ptr(dy), ld_dy, ptr(x), ld_x, rows, rows_dev, m, k, dw.data_ptr() + 4 * col0, dw.shape[1], scale,
                         None if db is None else db.data_ptr()))

    def _wgrad_flush(self):
        """Launches the queued weight gradients: on the side stream when they overlap the backward pass, the 64 x 64
        products of one row set as ONE tensor-core launch in TF32 mode (csrc/train_wgrad_tc.cu), the rest (and
        everything in fp32 mode, whose gradients the parity tests pin to fp32 FFMA sums) with the split-row fp32 kernel."""
        q, self._wq = self._wq, []
        if not q:
            return
        tc, rest = {}, []
        for w in q:
            dy, ld_dy, x, ld_x, rows, rows_dev, m, k = w[:8]
            if self.tc_wgrad and m == 64 and k == 64 and ld_dy % 4 == 0 and ld_x % 4 == 0 and (dy | x) % 16 == 0:
                tc.setdefault((rows, rows_dev), []).append(w)
            else:
                rest.append(w)

        def launch():
            for (rows, rows_dev), ws in tc.items():
                for i0 in range(0, len(ws), 6):
                    b = WgradBatch()
                    part = ws[i0:i0 + 6]
                    b.n_items, b.split, b.max_rows, b.n_rows_dev = len(part), 0, rows, rows_dev
                    b.workspace, b.workspace_bytes = self.ws_w.data_ptr(), self.ws_w.numel()
                    for it, w in zip(b.item, part):
                        it.dy, it.ld_dy, it.x, it.ld_x = w[0], w[1], w[2], w[3]
                        it.dw, it.ld_dw, it.scale, it.db = w[8], w[9], w[10], w[11]
                    L.check(self.lib.sgnn_linear_wgrad_tc(C.byref(b), self._stream()), launches=2)
            for (dy, ld_dy, x, ld_x, rows, rows_dev, m, k, dw, ld_dw, scale, db) in rest:
                L.check(self.lib.sgnn_linear_wgrad_ex(dy, ld_dy, x, ld_x, rows, rows_dev, m, k, dw, ld_dw, scale, db,
                                                      self.ws_w.data_ptr(), self.ws_w.numel(), self._stream()), launches=2)

        if not self.overlap_wgrad or self.profile is not None:
            with self._timed("wgrad_edges" if q[0][5] is not None else "wgrad_nodes"):
                launch()
            return
        ready = torch.cuda.Event()
        ready.record(torch.cuda.current_stream())          # the operands have been launched on the main stream
        self.side.wait_event(ready)
        with torch.cuda.stream(self.side):
            launch()

    def _ln_bwd(self, dy, index, act, ln, rows, rows_dev, dx):
        with self._timed("ln_bwd_edges" if rows_dev is not None else "ln_bwd_nodes"):
          L.check(self.lib.sgnn_layer_norm_bwd_rows(dy.data_ptr(), index, act["y2"].data_ptr(), act["mean"].data_ptr(),
                                                  act["rstd"].data_ptr(), ln.weight.data_ptr(), rows, rows_dev, self.h,
                                                  dx.data_ptr(), ln.weight.grad.data_ptr(), ln.bias.grad.data_ptr(),
                                                  self.ws.data_ptr(), self.ws.numel(), self._stream()), launches=2)

    def _segsum(self, m, rowptr, perm, out, ld_out):
        with self._timed("segment_sum"):
          L.check(self.lib.sgnn_segment_sum_rows_ld(m.data_ptr(), rowptr.data_ptr(), None if perm is None else perm.data_ptr(),
                                                  self.n, self.h, out, ld_out, self._stream()))

    # ------------------------------------------------------------------ the step
    def _forward_backward(self):
        n, h, cap, dim = self.n, self.h, self.cap, self.dim
        csr = self.rg.csr
        ed = self.n_edges_dev
        st = self.sim._stats(self.dev)
        win = self.window
        s = win.shape[1]
        with self._timed("pack_weights"):
            L.check(self.lib.sgnn_train_pack(self.descs.data_ptr(), self.n_descs, self.img_hi.data_ptr(),
                                             self.img_lo.data_ptr(), self._stream()))
        # noisy window, target, neighbour search on the noisy last frame (learned_simulator.py:461-470)
        torch.add(self.pos_in, self.noise, out=win)
        last = win[0, s - 1]
        with self._timed("neighbour_search"):
            self.rg.build(last, win.stride(0), dim, self.example_ptr, self.radius)
        # edges sorted by sender (stable): the backward of the sender gather is a segmented sum in edge order
        with self._timed("group_by_sender"):
            L.check(self.lib.sgnn_csr_by_source(csr.src.data_ptr(), ed, cap, n, self.rp_src.data_ptr(),
                                                self.perm_src.data_ptr(), self.ws_src.data_ptr(), self.ws_src.numel(),
                                                self._stream()), launches=5)
        L.check(self.lib.sgnn_node_features(win.data_ptr(), win.stride(0), n, s, dim, st["velocity"]["mean"].data_ptr(),
                                            st["velocity"]["std"].data_ptr(), 2.0, self.radius, 1.0, self.nf.data_ptr(),
                                            self.k_node_pad, self._stream()))
        # ---- forward ----
        m = self.mlp["ne"]
        hi, lo = self._img(m["w1"]); hi2, lo2 = self._img(m["w2"])
        a = self.act_ne
        self._fwd(tag="fwd_node_encoder", max_rows=n, prologue=TPRO_ROWS, a0=self.nf, ld0=self.k_node_pad, k0=self.k_node_pad, kpad=self.k_node_pad,
                  hidden=h, out=h, n_linear=2, has_ln=1, w1_hi=hi, w1_lo=lo, w2_hi=hi2, w2_lo=lo2, b1=m["a"].bias,
                  b1_cols=h, b2=m["b"].bias, gamma=m["ln"].weight, beta=m["ln"].bias, h1=a["h1"], y2=a["y2"],
                  mean=a["mean"], rstd=a["rstd"], y=self.x[0], ld_y=h)
        m = self.mlp["ee"]
        hi, lo = self._img(m["w1"]); hi2, lo2 = self._img(m["w2"])
        a = self.act_ee
        self._fwd(tag="fwd_edge_encoder", max_rows=cap, n_rows_dev=ed, prologue=TPRO_EDGE_FEATURES, pos=last.data_ptr(), pos_stride=win.stride(0),
                  dim=dim, radius=self.radius, src=csr.src, tgt=csr.tgt, feat_out=self.ef, kpad=32, hidden=h, out=h,
                  n_linear=2, has_ln=1, w1_hi=hi, w1_lo=lo, w2_hi=hi2, w2_lo=lo2, b1=m["a"].bias, b1_cols=h,
                  b2=m["b"].bias, gamma=m["ln"].weight, beta=m["ln"].bias, h1=a["h1"], y2=a["y2"], mean=a["mean"],
                  rstd=a["rstd"], y=self.e0, ld_y=h)
        for l in range(len(self.layers)):
            me, mn, ae, an = self.mlp_e[l], self.mlp_n[l], self.act_e[l], self.act_n[l]
            hi, lo = self._img(me["pq"])
            self._fwd(tag="fwd_pq", max_rows=n, prologue=TPRO_ROWS, a0=self.x[l], ld0=h, k0=h, kpad=h, hidden=2 * h, out=2 * h,
                      n_linear=1, w1_hi=hi, w1_lo=lo, b1=me["a"].bias, b1_cols=h, y=self.pq[l], ld_y=2 * h)
            hi, lo = self._img(me["w1"]); hi2, lo2 = self._img(me["w2"])
            self._fwd(tag="fwd_edge_mlp", max_rows=cap, n_rows_dev=ed, prologue=TPRO_ROWS, a0=self.e0, ld0=h, k0=h, kpad=h, src=csr.src,
                      tgt=csr.tgt, add_p=self.pq[l], add_q=self.pq[l].data_ptr() + 4 * h, ld_add=2 * h, hidden=h, out=h,
                      n_linear=2, has_ln=1, w1_hi=hi, w1_lo=lo, w2_hi=hi2, w2_lo=lo2, b2=me["b"].bias,
                      gamma=me["ln"].weight, beta=me["ln"].bias, h1=ae["h1"], y2=ae["y2"], mean=ae["mean"],
                      rstd=ae["rstd"], y=self.msg, ld_y=h)
            self._segsum(self.msg, csr.rowptr, None, self.agg[l].data_ptr(), h)
            hi, lo = self._img(mn["w1"]); hi2, lo2 = self._img(mn["w2"])
            self._fwd(tag="fwd_node_mlp", max_rows=n, prologue=TPRO_ROWS2, a0=self.agg[l], ld0=h, k0=h, a1=self.x[l], ld1=h, k1=h,
                      kpad=2 * h, hidden=h, out=h, n_linear=2, has_ln=1, w1_hi=hi, w1_lo=lo, w2_hi=hi2, w2_lo=lo2,
                      b1=mn["a"].bias, b1_cols=h, b2=mn["b"].bias, gamma=mn["ln"].weight, beta=mn["ln"].bias,
                      res=self.x[l], ld_res=h, h1=an["h1"], y2=an["y2"], mean=an["mean"], rstd=an["rstd"],
                      y=self.x[l + 1], ld_y=h)
        md = self.mlp["dec"]
        xl = self.x[-1]
        hi, lo = self._img(md["w1"]); hi2, lo2 = self._img(md["w2"])
        self._fwd(tag="fwd_decoder", max_rows=n, prologue=TPRO_ROWS, a0=xl, ld0=h, k0=h, kpad=h, hidden=h, out=dim + 1, n_linear=2,
                  has_ln=0, w1_hi=hi, w1_lo=lo, w2_hi=hi2, w2_lo=lo2, b1=md["a"].bias, b1_cols=h, b2=md["b"].bias,
                  h1=self.h1_d, y=self.pred, ld_y=4)

        # ---- loss (train.py:257-268) and its gradient ----
        acc_st = st["acceleration"]
        next_adj = self.next_pos + self.noise[:, -1]
        target = ((next_adj - win[:, -1]) - (win[:, -1] - win[:, -2]) - acc_st["mean"]) / acc_st["std"]
        d_acc = self.pred[:, :dim] - target
        d_str = self.pred[:, dim] - self.next_strain
        self.loss.copy_(((self.w_pos * (d_acc * d_acc).sum(-1) + self.w_strain * d_str * d_str).sum() * self.inv_n))
        self.dpred[:, :dim] = d_acc * (2.0 * self.w_pos * self.inv_n)
        self.dpred[:, dim] = d_str * (2.0 * self.w_strain * self.inv_n)

        # ---- backward ----
        dx, dx_new = self.dx
        nl = len(self.layers)
        self.dpre_n = self.dpre_n_all[nl + 1]
        hi, lo = self._img(md["w2t"]); hi1, lo1 = self._img(md["w1t"])
        self._bwd(tag="bwd_decoder", max_rows=n, dy=self.dpred, ld_dy=4, k_dy=4, n_linear=2, hidden=h, h1=self.h1_d, dpre=self.dpre_n,
                  w2t_hi=hi, w2t_lo=lo, want_dx=1, kx=h, w1t_hi=hi1, w1t_lo=lo1, dx0=dx, ld_dx0=h, k0=h)
        self._wgrad(self.dpred, 4, self.h1_d, h, n, None, dim + 1, h, md["b"].weight.grad, 0, db=md["b"].bias.grad)
        self._wgrad(self.dpre_n, h, xl, h, n, None, h, h, md["a"].weight.grad, 0, db=md["a"].bias.grad)
        self._wgrad_flush()
        self.de0.zero_()
        for l in reversed(range(len(self.layers))):
            me, mn, ae, an = self.mlp_e[l], self.mlp_n[l], self.act_e[l], self.act_n[l]
            self.dy2_n, self.dpre_n, self.dy2_e, self.dpre_e, self.dpq = (self.dy2_n_all[l], self.dpre_n_all[l], self.dy2_e_all[l],
                                                                           self.dpre_e_all[l], self.dpq_all[l])
            # node update: x' = x + LN(MLP([agg | x]))
            self._ln_bwd(dx, None, an, mn["ln"], n, None, self.dy2_n)
            hi, lo = self._img(mn["w2t"]); hi1, lo1 = self._img(mn["w1t"])
            self._bwd(tag="bwd_node_mlp", max_rows=n, dy=self.dy2_n, ld_dy=h, k_dy=h, n_linear=2, hidden=h, h1=an["h1"], dpre=self.dpre_n,
                      w2t_hi=hi, w2t_lo=lo, want_dx=1, kx=2 * h, w1t_hi=hi1, w1t_lo=lo1, dx0=self.dagg, ld_dx0=h, k0=h,
                      dx1=dx_new, ld_dx1=h, k1=h, add1=dx, ld_add1=h)
            self._wgrad(self.dy2_n, h, an["h1"], h, n, None, h, h, mn["b"].weight.grad, 0, db=mn["b"].bias.grad)
            self._wgrad(self.dpre_n, h, self.agg[l], h, n, None, h, h, mn["a"].weight.grad, 0, db=mn["a"].bias.grad)
            self._wgrad(self.dpre_n, h, self.x[l], h, n, None, h, h, mn["a"].weight.grad, h)
            # messages: agg = segment sum of LN(MLP(...)); its backward gathers dagg by receiver
            self._ln_bwd(self.dagg, csr.tgt.data_ptr(), ae, me["ln"], cap, ed, self.dy2_e)
            hi, lo = self._img(me["w2t"]); hi1, lo1 = self._img(me["w1t"])
            self._bwd(tag="bwd_edge_mlp", max_rows=cap, n_rows_dev=ed, dy=self.dy2_e, ld_dy=h, k_dy=h, n_linear=2, hidden=h, h1=ae["h1"],
                      dpre=self.dpre_e, w2t_hi=hi, w2t_lo=lo, want_dx=1, kx=h, w1t_hi=hi1, w1t_lo=lo1, dx0=self.de0,
                      ld_dx0=h, k0=h, acc0=1)
            self._wgrad(self.dy2_e, h, ae["h1"], h, cap, ed, h, h, me["b"].weight.grad, 0, db=me["b"].bias.grad)
            self._wgrad(self.dpre_e, h, self.e0, h, cap, ed, h, h, me["a"].weight.grad, 2 * h, scale=me["scale"],
                        db=me["a"].bias.grad)
            self._wgrad_flush()
            self._segsum(self.dpre_e, csr.rowptr, None, self.dpq.data_ptr(), 2 * h)
            self._segsum(self.dpre_e, self.rp_src, self.perm_src, self.dpq.data_ptr() + 4 * h, 2 * h)
            hi1, lo1 = self._img(me["pqt"])
            self._bwd(tag="bwd_pq", max_rows=n, dy=self.dpq, ld_dy=2 * h, k_dy=2 * h, n_linear=1, want_dx=1, kx=h, w1t_hi=hi1,
                      w1t_lo=lo1, dx0=dx_new, ld_dx0=h, k0=h, acc0=1)
            self._wgrad(self.dpq, 2 * h, self.x[l], h, n, None, h, h, me["a"].weight.grad, 0)
            self._wgrad(self.dpq.data_ptr() + 4 * h, 2 * h, self.x[l], h, n, None, h, h, me["a"].weight.grad, h)
            self._wgrad_flush()
            dx, dx_new = dx_new, dx
        # encoders
        self.dy2_n, self.dpre_n, self.dy2_e, self.dpre_e = (self.dy2_n_all[nl], self.dpre_n_all[nl], self.dy2_e_all[nl],
                                                            self.dpre_e_all[nl])
        m, a = self.mlp["ee"], self.act_ee
        self._ln_bwd(self.de0, None, a, m["ln"], cap, ed, self.dy2_e)
        hi, lo = self._img(m["w2t"])
        self._bwd(tag="bwd_edge_encoder", max_rows=cap, n_rows_dev=ed, dy=self.dy2_e, ld_dy=h, k_dy=h, n_linear=2, hidden=h, h1=a["h1"],
                  dpre=self.dpre_e, w2t_hi=hi, w2t_lo=lo, want_dx=0)
        self._wgrad(self.dy2_e, h, a["h1"], h, cap, ed, h, h, m["b"].weight.grad, 0, db=m["b"].bias.grad)
        self._wgrad(self.dpre_e, h, self.ef, dim + 1, cap, ed, h, dim + 1, m["a"].weight.grad, 0, db=m["a"].bias.grad)
        self._wgrad_flush()
        m, a = self.mlp["ne"], self.act_ne
        self._ln_bwd(dx, None, a, m["ln"], n, None, self.dy2_n)
        hi, lo = self._img(m["w2t"])
        self._bwd(tag="bwd_node_encoder", max_rows=n, dy=self.dy2_n, ld_dy=h, k_dy=h, n_linear=2, hidden=h, h1=a["h1"], dpre=self.dpre_n,
                  w2t_hi=hi, w2t_lo=lo, want_dx=0)
        self._wgrad(self.dy2_n, h, a["h1"], h, n, None, h, h, m["b"].weight.grad, 0, db=m["b"].bias.grad)
        self._wgrad(self.dpre_n, h, self.nf, self.k_node_pad, n, None, h, self.k_node, m["a"].weight.grad, 0, db=m["a"].bias.grad)
        self._wgrad_flush()
        if self.overlap_wgrad and self.profile is None:         # every gradient is in the flat buffer before allreduce / Adam
            done = torch.cuda.Event()
            done.record(self.side)
            torch.cuda.current_stream().wait_event(done)

    def _whole_step(self):
        with self._timed("forward_backward_total"):
            self._forward_backward()
        if self.world > 1:
            with self._timed("allreduce"):
                dist.all_reduce(self.flat_grad, op=dist.ReduceOp.SUM, group=self.group)
        with self._timed("adam"):
            self.opt.step()

    # ------------------------------------------------------------------ public
    def _bind(self, batch: Dict[str, torch.Tensor]):
        pos = batch["positions"]
        n, s, dim = pos.shape
        if n != self.n or dim != self.dim:
            raise L.SgnnError(f"TrainEngine built for {self.n} particles of dimension {self.dim}, got {n} x {dim}")
        if self.window is None or self.window.shape[1] != s:
            self.window = torch.zeros(n, s, dim, dtype=torch.float32, device=self.dev)
            self.pos_in = torch.zeros_like(self.window)
            self.noise = torch.zeros_like(self.window)
            self.graph = None
        self.pos_in.copy_(pos)
        self.next_pos.copy_(batch["next_position"])
        self.next_strain.copy_(batch["next_strain"])
        ptr = example_ptr_of(batch["n_particles_per_example"], self.dev)
        if self.example_ptr is None or self.example_ptr.numel() != ptr.numel():
            self.example_ptr = ptr.clone()
            self.graph = None
        else:
            self.example_ptr.copy_(ptr)

    def step(self, batch: Dict[str, torch.Tensor], noise_std: float = 0.0, noise: Optional[torch.Tensor] = None,
             generator: Optional[torch.Generator] = None, n_global: Optional[int] = None) -> torch.Tensor:
        """One optimisation step.  batch: the tensors of the reference's collate_fn, on the device.  Returns this
        rank's contribution to the global mean loss as a device scalar (sum over ranks = the loss)."""
        L.require_cuda(batch["positions"], "positions")
        if tuple(p.data_ptr() for p in self.params) != self._param_ptrs:
            raise L.SgnnError("TrainEngine: the model's parameters were re-allocated; build a new engine")
        self._bind(batch)
        if noise is None:
            noise = noise_utils.get_random_walk_noise_for_position_sequence(self.pos_in, noise_std, generator)
        self.noise.copy_(noise)
        self.inv_n.fill_(1.0 / float(self.n * self.world if n_global is None else n_global))
        self._attach_grads()
        if not self.use_graph:
            self._whole_step()
            return self.loss[0].clone()
        if self.graph is None:
            side = torch.cuda.Stream(device=self.dev)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):                   # shared-memory attributes, allocator pools, NCCL set-up
                self._snapshot()
                self._whole_step()
                self._restore()
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self._whole_step()
            self._restore()                                 # the capture itself executes nothing
        self.graph.replay()
        return self.loss[0].clone()

    def forward_backward(self, batch: Dict[str, torch.Tensor], noise: torch.Tensor, n_global: Optional[int] = None
                         ) -> torch.Tensor:
        """Forward + backward only, un-captured (tests): fills every p.grad, returns the loss."""
        self._bind(batch)
        self.noise.copy_(noise)
        self.inv_n.fill_(1.0 / float(self.n * self.world if n_global is None else n_global))
        self._attach_grads()
        self._forward_backward()
        return self.loss[0].clone()

    # the warm-up step before a capture must not count as a training step
    def _snapshot(self):
        self._saved_params = [p.detach().clone() for p in self.params]
        self._saved_state = {id(p): {k: (v.clone() if isinstance(v, torch.Tensor) else v) for k, v in self.opt.state.get(p, {}).items()}
                             for p in self.params}

    def _restore(self):
        with torch.no_grad():
            for p, q in zip(self.params, self._saved_params):
                p.copy_(q)
            for p in self.params:
                cur, old = self.opt.state.get(p, {}), self._saved_state[id(p)]
                for k, v in cur.items():
                    if isinstance(v, torch.Tensor):
                        if k in old:
                            v.copy_(old[k])
                        else:
                            v.zero_()
