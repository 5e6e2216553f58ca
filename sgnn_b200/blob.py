"""Builds the per-kernel weight blobs (see csrc/params.h): GEMM weights as K-major SWIZZLE_128B
images (packed on the device by sgnn_pack_weight), followed by the fp32 vectors."""
import ctypes as C
from typing import List, Optional, Tuple

import torch

from . import _lib as L


VEC_REGION_BYTES = 8192     # SGNN_VEC_REGION_BYTES in csrc/params.h


def pad32(k: int) -> int:
    return (k + 31) // 32 * 32


class BlobBuilder:
    """Two passes: plan offsets (add_image / add_vector), then `finish()` allocates the device
    buffer and fills it.  Image offsets are bytes (1024-aligned), vector offsets are floats."""

    def __init__(self, device, round_tf32: bool):
        self.device = device
        self.round_tf32 = round_tf32
        self._images = []   # (byte_off, weight, row0, n, col0, k, kpad, scale)
        self._vectors = []  # (float_off, tensor)
        self._bytes = VEC_REGION_BYTES     # images start after the vector region
        self._vec_bytes = 0

    def add_image(self, weight: torch.Tensor, row0: int, n: int, col0: int, k: int,
                  kpad: Optional[int] = None, scale: float = 1.0) -> int:
        """weight[row0:row0+n, col0:col0+k] of an nn.Linear weight ([out, in]); n padded to 16."""
        kpad = pad32(k) if kpad is None else kpad
        n_pad = (n + 15) // 16 * 16
        assert n_pad == n, "GEMM output widths must be multiples of 16"
        off = (self._bytes + 1023) // 1024 * 1024
        self._images.append((off, weight, row0, n, col0, k, kpad, float(scale)))
        self._bytes = off + n * kpad * 4
        return off

    def add_vector(self, t: torch.Tensor) -> int:
        off = (self._vec_bytes + 15) // 16 * 16
        if off + t.numel() * 4 > VEC_REGION_BYTES:
            raise L.SgnnError(f"fp32 vectors of this kernel exceed the {VEC_REGION_BYTES}-byte blob region")
        self._vectors.append((off // 4, t))
        self._vec_bytes = off + t.numel() * 4
        return off // 4

    def finish(self) -> Tuple[torch.Tensor, int]:
        total = (self._bytes + 15) // 16 * 16
        blob = torch.zeros(total // 4, dtype=torch.float32, device=self.device)
        lib = L.lib()
        keep = []
        for off, w, row0, n, col0, k, kpad, scale in self._images:
            w = w.detach()
            if w.dtype != torch.float32 or not w.is_contiguous():
                w = w.float().contiguous()
            keep.append(w)
            L.check(lib.sgnn_pack_weight(L.ptr(w), w.shape[1], row0, n, col0, k, kpad,
                                         C.c_float(scale), int(self.round_tf32),
                                         C.c_void_p(blob.data_ptr() + off), L.stream_ptr()))
        for foff, t in self._vectors:
            blob[foff:foff + t.numel()].copy_(t.detach().reshape(-1).float())
        # kernels stage the blob BEFORE their programmatic-dependent-launch wait (it is constant
        # for them): make sure the packing above has finished before anything can be launched
        torch.cuda.current_stream().synchronize()
        return blob, total


def linears_of(seq: torch.nn.Module) -> List[torch.nn.Linear]:
    """The nn.Linear layers "NN-0", "NN-1", ... of a build_mlp Sequential, in order."""
    return [m for m in seq.children() if isinstance(m, torch.nn.Linear)]


def fill_mlp(b: BlobBuilder, mlp: L.Mlp, linears: List[torch.nn.Linear],
             ln: Optional[torch.nn.LayerNorm], first_as_image: bool = True, first_cols=None,
             first_kpad: Optional[int] = None, first_scale: float = 1.0, first_bias: bool = True,
             last_as_image: bool = True):
    """Describe Linear-ReLU-...-Linear[-LN] in `mlp` and register its tensors with `b`.
    first_cols=(c0, k): use only columns [c0, c0+k) of the first Linear (split concat weights)."""
    n_lin = len(linears)
    assert 2 <= n_lin <= L.MAX_HIDDEN + 1
    mlp.n_hidden = n_lin - 1
    mlp.hidden = linears[0].out_features
    mlp.out = linears[-1].out_features
    mlp.has_ln = int(ln is not None)
    for i, lin in enumerate(linears):
        w = lin.weight
        is_first, is_last = i == 0, i == n_lin - 1
        if is_first:
            c0, k = first_cols if first_cols is not None else (0, w.shape[1])
            mlp.w_off[i] = b.add_image(w, 0, w.shape[0], c0, k, first_kpad, first_scale) \
                if first_as_image else -1
            mlp.b_off[i] = b.add_vector(lin.bias) if first_bias else -1
        elif is_last and not last_as_image:
            mlp.w_off[i] = -1
            mlp.b_off[i] = b.add_vector(lin.bias)
        else:
            mlp.w_off[i] = b.add_image(w, 0, w.shape[0], 0, w.shape[1])
            mlp.b_off[i] = b.add_vector(lin.bias)
    if ln is not None:
        mlp.gamma_off = b.add_vector(ln.weight)
        mlp.beta_off = b.add_vector(ln.bias)
    else:
        mlp.gamma_off = mlp.beta_off = -1


def params_version(params) -> Tuple:
    """Cheap fingerprint to detect optimizer updates / load_state_dict / .to()."""
    return tuple((p.data_ptr(), p._version) for p in params)
