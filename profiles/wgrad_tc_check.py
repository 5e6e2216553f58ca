"""sgnn_linear_wgrad_tc against an fp64 product and the fp32 FFMA kernel: errors and time per call.
python profiles/wgrad_tc_check.py [rows]"""
import ctypes as C
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from sgnn_b200 import _lib as L  # noqa: E402
from sgnn_b200.train_engine import WgradBatch  # noqa: E402


def run(lib, items, rows, n_dev, split, ws):
    b = WgradBatch()
    b.n_items, b.split, b.max_rows = len(items), split, rows
    b.n_rows_dev = None if n_dev is None else n_dev.data_ptr()
    b.workspace, b.workspace_bytes = ws.data_ptr(), ws.numel()
    for i, (dy, x, dw, db, scale) in enumerate(items):
        it = b.item[i]
        it.dy, it.x, it.dw, it.db = dy.data_ptr(), x.data_ptr(), dw.data_ptr(), None if db is None else db.data_ptr()
        it.ld_dy, it.ld_x, it.ld_dw, it.scale = dy.stride(0), x.stride(0), dw.stride(0), scale
    L.check(lib.sgnn_linear_wgrad_tc(C.byref(b), L.stream_ptr()), 2)


if __name__ == "__main__":
    rows = int(sys.argv[1]) if len(sys.argv) > 1 else 172800
    lib = L.lib()
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    cap = rows + 1000
    live = torch.tensor([rows], dtype=torch.int32, device=dev)
    ws = torch.empty(lib.sgnn_linear_wgrad_tc_workspace_bytes(1, cap), dtype=torch.uint8, device=dev)
    wide = torch.randn(cap, 128, device=dev)
    dy, x = wide[:, 64:], torch.randn(cap, 64, device=dev)         # dy: a column block (ld 128)
    ref = (dy[:rows].double().T @ x[:rows].double())
    refb = dy[:rows].double().sum(0)
    for split in (0, 1):
        dw = torch.zeros(64, 192, device=dev)
        db = torch.zeros(64, device=dev)
        run(lib, [(dy, x, dw[:, 64:128], db, 0.5)], cap, live, split, ws)
        torch.cuda.synchronize()
        got = dw[:, 64:128].double() / 0.5
        scale = ref.abs().max().item()
        if os.environ.get("WGTC_DEBUG"):
            print("got", got[:2, :6].tolist(), "\nref", ref[:2, :6].tolist(), "\ndb", db[:4].tolist(), refb[:4].tolist())
            part = ws.view(torch.float32)[:4160]
            print("partial 0: nonzero", int((part != 0).sum()), part[:6].tolist())
        print(f"split={split} rows={rows}: max|dW - ref| / max|ref| = {(got - ref).abs().max().item() / scale:.3e}, "
              f"db {(db.double() - refb).abs().max().item() / refb.abs().max().item():.3e}, "
              f"outside block untouched: {bool((dw[:, :64] == 0).all() and (dw[:, 128:] == 0).all())}")
        dw2 = torch.zeros_like(dw)
        run(lib, [(dy, x, dw2[:, 64:128], db, 0.5)], cap, live, split, ws)
        print("  bitwise repeatable:", torch.equal(dw, dw2))
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(50):
            run(lib, [(dy, x, dw[:, 64:128], db, 0.5)], cap, live, split, ws)
        t1.record()
        torch.cuda.synchronize()
        print(f"  {t0.elapsed_time(t1) / 50 * 1e3:.1f} us per call (kernel + finish)")
    # the fp32 FFMA kernel on the same operands
    ws2 = torch.empty(lib.sgnn_linear_wgrad_workspace_bytes(cap, 64, 64), dtype=torch.uint8, device=dev)
    dw = torch.zeros(64, 192, device=dev)
    db = torch.zeros(64, device=dev)
    args = (dy.data_ptr(), 128, x.data_ptr(), 64, cap, live.data_ptr(), 64, 64, dw.data_ptr() + 4 * 64, 192, 0.5,
            db.data_ptr(), ws2.data_ptr(), ws2.numel(), L.stream_ptr())
    vp, lg, it_, fl, sz = C.c_void_p, C.c_long, C.c_int, C.c_float, C.c_size_t
    lib.sgnn_linear_wgrad_ex.argtypes = [vp, lg, vp, lg, lg, vp, it_, it_, vp, lg, fl, vp, vp, sz, vp]
    L.check(lib.sgnn_linear_wgrad_ex(*args), 2)
    torch.cuda.synchronize()
    got = dw[:, 64:128].double() / 0.5
    print(f"fp32 FFMA kernel: {(got - ref).abs().max().item() / ref.abs().max().item():.3e}")
    t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0.record()
    for _ in range(50):
        L.check(lib.sgnn_linear_wgrad_ex(*args), 2)
    t1.record()
    torch.cuda.synchronize()
    print(f"  {t0.elapsed_time(t1) / 50 * 1e3:.1f} us per call")
