"""TEST INFRASTRUCTURE ONLY -- Python front-end of the neighbour-search oracle.

Restates `torch_geometric.nn.radius_graph` -> `torch_cluster.radius` (CUDA rule: first K
sources in ascending index order with fp32 d^2 < (float)(r*r), same example only).  Called
by the reference at sgnn/single_scale/learned_simulator.py:116-117 and
sgnn/multi_scale/multi_scale_graph.py:222-227,269-274.  torch_cluster / torch_geometric are
third-party, not vendored and not pinned by the reference: parity for this piece is
"unpinned" (see radius_oracle.c header).

Only tests/, __graft_entry__.smoke() and bench.py's CPU-baseline legs may import this.
"""
import ctypes
import os
import subprocess
from concurrent.futures import ThreadPoolExecutor

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_ref", "libsgnn_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile radius_oracle.c into oracle/_ref/ (gcc, a few hundred ms)."""
    src = os.path.join(_HERE, "radius_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-s", "-C", _HERE])
    return _SO


def _load():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        _lib.sgnn_oracle_radius_graph.restype = ctypes.c_long
        _lib.sgnn_oracle_radius_graph_cells.restype = ctypes.c_long
    return _lib


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def _split(n, threads):
    threads = max(1, min(int(threads), (n + 255) // 256))
    step = (n + threads - 1) // threads
    return [(lo, min(n, lo + step)) for lo in range(0, n, step)]


def _run(fn, n, threads):
    """Run fn(q_lo, q_hi) over [0, n) on `threads` host threads (ctypes releases the GIL)."""
    chunks = _split(n, threads)
    if len(chunks) == 1:
        fn(*chunks[0])
    else:
        with ThreadPoolExecutor(len(chunks)) as ex:
            list(ex.map(lambda c: fn(*c), chunks))


def _compact(nbr, cnt, k):
    """[n, k+1] slab + counts -> edge_index [2, E] int64 (row 0 = source, row 1 = target)."""
    n = cnt.shape[0]
    mask = np.arange(k + 1)[None, :] < cnt[:, None]
    src = nbr[mask].astype(np.int64)
    tgt = np.repeat(np.arange(n, dtype=np.int64), cnt)
    return np.stack([src, tgt])


def radius_graph_np(pos, r, n_per_example=None, loop=True, max_num_neighbors=32,
                    use_fma=True, method="auto", threads=None):
    """numpy in / numpy out.  pos [n, dim] float32; returns edge_index [2, E] int64."""
    pos = np.ascontiguousarray(pos, dtype=np.float32)
    n, dim = pos.shape
    if n_per_example is None:
        n_per_example = [n]
    ptr = np.zeros(len(n_per_example) + 1, dtype=np.int64)
    ptr[1:] = np.cumsum(np.asarray(n_per_example, dtype=np.int64))
    assert ptr[-1] == n, "n_per_example must sum to the number of particles"
    k = int(max_num_neighbors)
    assert k + 1 <= 64
    nbr = np.zeros((n, k + 1), dtype=np.int32)
    cnt = np.zeros(n, dtype=np.int32)
    lib = _load()
    if threads is None:
        threads = os.cpu_count() or 1
    if method == "auto":
        method = "brute" if n <= 20000 else "cells"
    if n == 0:
        return np.zeros((2, 0), dtype=np.int64)
    if method == "brute":
        _run(lambda lo, hi: lib.sgnn_oracle_radius_graph(
            _ptr(pos), n, dim, _ptr(ptr), len(n_per_example), ctypes.c_double(float(r)), k,
            int(bool(loop)), int(bool(use_fma)), lo, hi, _ptr(nbr), _ptr(cnt)), n, threads)
    else:
        assert dim in (2, 3)
        # cell edge slightly above r so that fp rounding of the cell index can never hide a
        # true neighbour: |x_i - x_j| < r  =>  cell indices differ by at most 1.
        h = float(r) * (1.0 + 1e-5) + 1e-30
        lo = pos.min(axis=0).astype(np.float64)
        cxyz = np.zeros((n, 3), dtype=np.int32)
        cxyz[:, :dim] = np.floor((pos.astype(np.float64) - lo) / h).astype(np.int32)
        ncell = (cxyz.max(axis=0) + 1).astype(np.int32)
        example_of = (np.searchsorted(ptr, np.arange(n), side="right") - 1).astype(np.int32)
        nc = int(ncell[0]) * int(ncell[1]) * int(ncell[2])
        key = (example_of.astype(np.int64) * nc
               + (cxyz[:, 2].astype(np.int64) * ncell[1] + cxyz[:, 1]) * ncell[0] + cxyz[:, 0])
        order = np.argsort(key, kind="stable").astype(np.int32)
        counts = np.bincount(key, minlength=nc * len(n_per_example))
        cell_start = np.zeros(counts.shape[0] + 1, dtype=np.int64)
        cell_start[1:] = np.cumsum(counts)
        _run(lambda lo, hi: lib.sgnn_oracle_radius_graph_cells(
            _ptr(pos), n, dim, _ptr(example_of), _ptr(cxyz), _ptr(order), _ptr(cell_start),
            _ptr(ncell), ctypes.c_double(float(r)), k, int(bool(loop)), int(bool(use_fma)),
            lo, hi, _ptr(nbr), _ptr(cnt)), n, threads)
    return _compact(nbr, cnt, k)


def radius_graph_py(pos, r, n_per_example=None, loop=True, max_num_neighbors=32):
    """Pure-Python/numpy loop form of the same rule (small cases only; cross-checks the C)."""
    pos = np.asarray(pos, dtype=np.float32)
    n, dim = pos.shape
    if n_per_example is None:
        n_per_example = [n]
    r2 = np.float32(float(r) * float(r))
    kq = max_num_neighbors if loop else max_num_neighbors + 1
    src, tgt = [], []
    lo = 0
    for ne in n_per_example:
        hi = lo + int(ne)
        for i in range(lo, hi):
            c = 0
            for j in range(lo, hi):
                dist = np.float32(0.0)
                for d in range(dim):
                    diff = np.float32(pos[j, d] - pos[i, d])
                    # fma(diff, diff, dist): product exact in float64 (24+24 bits), one rounding
                    dist = np.float32(np.float64(diff) * np.float64(diff) + np.float64(dist))
                if dist < r2:
                    if loop or i != j:
                        src.append(j)
                        tgt.append(i)
                    c += 1
                    if c >= kq:
                        break
        lo = hi
    return np.array([src, tgt], dtype=np.int64).reshape(2, -1)
