"""Per-call time of the Linear weight-gradient kernel (csrc/train_linear.cu) against the library GEMM it
replaces (dy^T x through torch.mm + a column sum), at the shapes of the bench train step:
    python profiles/wgrad_micro.py"""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sgnn_b200 import _lib as L  # noqa: E402


def timed(fn, iters=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters * 1e3


def main():
    dev = torch.device("cuda", 0)
    lib = L.lib()
    print("rows m k | kernel us (GB/s, TFLOP/s) | torch mm + sum us")
    for rows, m, k in [(38400, 64, 64), (190000, 64, 64), (190000, 64, 3), (38400, 64, 30), (1000000, 64, 64)]:
        dy = torch.randn(rows, m, device=dev)
        x = torch.randn(rows, k, device=dev)
        dw = torch.empty(m, k, device=dev)
        db = torch.empty(m, device=dev)
        nbytes = lib.sgnn_linear_wgrad_workspace_bytes(C.c_long(rows), m, k)
        ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)

        def ours():
            L.check(lib.sgnn_linear_wgrad(L.ptr(dy), L.ptr(x), C.c_long(rows), m, k, L.ptr(dw), L.ptr(db), L.ptr(ws),
                                          C.c_size_t(nbytes), L.stream_ptr()), launches=2)

        def lib_gemm():
            return dy.t() @ x, dy.sum(0)

        t = timed(ours)
        t2 = timed(lib_gemm)
        want = dy.double().t() @ x.double()
        err = (dw.double() - want).abs().max().item() / want.abs().max().item()
        print(f"{rows} {m} {k} | {t:.1f} us ({4 * rows * (m + k) / t / 1e3:.0f} GB/s, {2 * rows * m * k / t / 1e6:.1f} TFLOP/s, "
              f"rel err {err:.1e}, {nbytes >> 10} KB partials) | {t2:.1f} us")


if __name__ == "__main__":
    main()
