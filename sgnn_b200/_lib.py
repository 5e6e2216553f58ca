"""ctypes binding of libsgnn_b200.so (the C ABI declared in include/sgnn_b200.h).

There is deliberately NO fallback: if the shared library is missing or a call fails, the
caller gets an exception.  PyTorch is used for device memory and streams only.
"""
import ctypes as C
import os

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libsgnn_b200.so")
MAX_HIDDEN = 3

PRO_FEATURES, PRO_AGG_X, PRO_ROWS = 0, 1, 2
TAIL_NONE, TAIL_PQ, TAIL_DECODE, TAIL_PX = 0, 1, 2, 3


class SgnnError(RuntimeError):
    pass


class Mlp(C.Structure):
    _fields_ = [("n_hidden", C.c_int), ("hidden", C.c_int), ("out", C.c_int), ("has_ln", C.c_int),
                ("w_off", C.c_int * (MAX_HIDDEN + 1)), ("b_off", C.c_int * (MAX_HIDDEN + 1)),
                ("gamma_off", C.c_int), ("beta_off", C.c_int)]


class EdgeLayerParams(C.Structure):
    _fields_ = [("e_tiles", C.c_void_p), ("pq", C.c_void_p), ("src", C.c_void_p),
                ("tgt", C.c_void_p), ("rowptr", C.c_void_p), ("n_nodes", C.c_int), ("he", C.c_int),
                ("blob", C.c_void_p), ("blob_bytes", C.c_int), ("mlp", Mlp), ("agg", C.c_void_p),
                ("head", C.c_void_p), ("max_tiles", C.c_int), ("wxs_off", C.c_int), ("defer_affine", C.c_int)]


class EdgeEncoderParams(C.Structure):
    _fields_ = [("pos", C.c_void_p), ("pos_stride", C.c_long), ("dim", C.c_int),
                ("inv_radius_unused", C.c_float), ("radius", C.c_float), ("src", C.c_void_p),
                ("tgt", C.c_void_p), ("rowptr", C.c_void_p), ("n_nodes", C.c_int),
                ("feat", C.c_void_p), ("k_in", C.c_int), ("blob", C.c_void_p),
                ("blob_bytes", C.c_int), ("mlp", Mlp), ("w0_off", C.c_int), ("e_tiles", C.c_void_p),
                ("feat_out", C.c_void_p), ("max_tiles", C.c_int)]


class NodeParams(C.Structure):
    _fields_ = [("n_rows", C.c_int), ("prologue", C.c_int),
                ("pos_seq", C.c_void_p), ("pos_seq_stride", C.c_long), ("seq_len", C.c_int), ("dim", C.c_int),
                ("vel_mean", C.c_void_p), ("vel_std", C.c_void_p), ("wall_offset", C.c_float),
                ("wall_max", C.c_float), ("wall_div", C.c_float), ("types", C.c_void_p),
                ("emb", C.c_void_p), ("emb_dim", C.c_int), ("feat_out", C.c_void_p),
                ("agg", C.c_void_p), ("head", C.c_void_p), ("rowptr", C.c_void_p),
                ("agg_w", C.c_int), ("x_in", C.c_void_p), ("x_w", C.c_int),
                ("rows_in", C.c_void_p), ("rows_ld", C.c_long), ("k_in", C.c_int),
                ("k_pad", C.c_int), ("blob", C.c_void_p), ("blob_bytes", C.c_int),
                ("has_mlp", C.c_int), ("mlp", Mlp), ("residual", C.c_int), ("x_out", C.c_void_p),
                ("tail", C.c_int), ("pq_hidden", C.c_int), ("wp_off", C.c_int), ("wq_off", C.c_int),
                ("pq_bias_off", C.c_int), ("pq_out", C.c_void_p), ("dec", Mlp),
                ("dec_wlast_off", C.c_int), ("dec_out", C.c_void_p), ("acc_mean", C.c_void_p),
                ("acc_std", C.c_void_p), ("next_pos", C.c_void_p), ("next_pos_stride", C.c_long),
                ("strain_out", C.c_void_p), ("xr_in", C.c_void_p), ("xr_ld", C.c_long),
                ("n_types", C.c_int), ("agg_affine", C.c_int), ("agg_gamma_off", C.c_int), ("agg_beta_off", C.c_int)]


_lib = None
launch_count = 0          # kernels launched through this binding (bench.py reports it)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise SgnnError(
                f"{LIB_PATH} not found: build it with `python sgnn_b200/build_lib.py` "
                "(there is no CPU or PyTorch fallback for the sgnn_b200 kernels)")
        _lib = C.CDLL(LIB_PATH)
        _lib.sgnn_last_error.restype = C.c_char_p
        _lib.sgnn_radius_graph_workspace_bytes.restype = C.c_size_t
        _lib.sgnn_edge_layer_smem_bytes.restype = C.c_size_t
        _lib.sgnn_edge_encoder_smem_bytes.restype = C.c_size_t
        _lib.sgnn_node_smem_bytes.restype = C.c_size_t
        _lib.sgnn_layer_norm_bwd_workspace_bytes.restype = C.c_size_t
        _lib.sgnn_linear_wgrad_workspace_bytes.restype = C.c_size_t
        _lib.sgnn_csr_by_source_workspace_bytes.restype = C.c_size_t
        _lib.sgnn_linear_wgrad_tc_workspace_bytes.restype = C.c_size_t
        _lib.sgnn_linear_wgrad_tc_workspace_bytes.argtypes = [C.c_int, C.c_long]
        # training-path entries are called ~100 times per step from Python: declared argument types spare the
        # per-call type inference of ctypes (the call sites keep passing c_void_p / c_long / ints / None)
        vp, lg, it, fl, sz = C.c_void_p, C.c_long, C.c_int, C.c_float, C.c_size_t
        _lib.sgnn_gather_add_rows.argtypes = [vp, lg, vp, lg, vp, vp, it, it, vp, vp]
        _lib.sgnn_gather_add_rows_acc.argtypes = [vp, lg, vp, lg, vp, vp, it, it, vp, vp]
        _lib.sgnn_segment_sum_rows.argtypes = [vp, vp, vp, it, it, vp, vp]
        _lib.sgnn_segment_sum_rows_ld.argtypes = [vp, vp, vp, it, it, vp, lg, vp]
        _lib.sgnn_gather_rows.argtypes = [vp, lg, vp, it, it, vp, vp]
        _lib.sgnn_layer_norm_fwd.argtypes = [vp, lg, it, vp, vp, fl, vp, vp, vp, vp]
        _lib.sgnn_layer_norm_bwd.argtypes = [vp, vp, vp, vp, vp, lg, it, vp, vp, vp, vp, sz, vp]
        _lib.sgnn_layer_norm_bwd_static.argtypes = [vp, vp, vp, vp, vp, lg, it, vp, vp, vp, vp]
    return _lib


def check(rc, launches=1):
    global launch_count
    if rc != 0:
        raise SgnnError(lib().sgnn_last_error().decode() or f"libsgnn_b200 error {rc}")
    launch_count += launches


def stream_ptr():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    if t is None:
        return None
    return C.c_void_p(t.data_ptr())


_dev_info = {}


def device_info():
    """(sm_count, cc_major, cc_minor) of the current CUDA device; raises off sm_100."""
    dev = torch.cuda.current_device()
    if dev not in _dev_info:
        sm, ma, mi = C.c_int(), C.c_int(), C.c_int()
        check(lib().sgnn_device_info(C.byref(sm), C.byref(ma), C.byref(mi)), 0)
        if ma.value != 10:
            raise SgnnError(f"sgnn_b200 kernels are built for sm_100a only; device has sm_{ma.value}{mi.value}")
        _dev_info[dev] = (sm.value, ma.value, mi.value)
    return _dev_info[dev]


def require_cuda(t, name):
    if not t.is_cuda:
        raise SgnnError(f"{name} must be a CUDA tensor: sgnn_b200 has no CPU path")
