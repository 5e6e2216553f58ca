"""In-kernel timeline of edge_layer_ws_kernel (CTA 0) at ~1M particles.  Needs a library built with
SGNN_TIMELINE=1 (python -c "from sgnn_b200 import build_lib; build_lib.build(force=True)")."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from sgnn_b200 import _lib as L  # noqa: E402
from sgnn_b200.single_scale import evaluate as EV  # noqa: E402

dev = torch.device("cuda", 0)
nx, ny = 2048, 512
traj, stats, radius = bench.make_case(seed=7, nx=nx, ny=ny, nsteps=2)
sim = bench.make_sim(stats, radius, dev)
n = nx * ny
plan = EV.RolloutPlan(sim, n, bench.S + 2, bench.S, 2, "autoregressive", dev, use_graph=False)
plan.traj[:, :bench.S].copy_(traj["positions"][:, :bench.S].to(dev))
plan.launch_steps(0, 1)
torch.cuda.synchronize()
for flags in [int(a) for a in (sys.argv[1:] or ["0"])]:
    L.check(L.lib().sgnn_debug_ws_flags(flags))
    buf = torch.zeros(256 * 16, dtype=torch.int64, device=dev)
    L.check(L.lib().sgnn_debug_ws_timeline(C.c_void_p(buf.data_ptr())))
    ms, e = bench.measure_edge_layer(sim, n, dev, reps=1)
    torch.cuda.synchronize()
    L.check(L.lib().sgnn_debug_ws_timeline(None))
    t = buf.cpu().view(256, 16)
    t0 = int(t[0, 0])
    print(f"flags {flags}: {ms:.3f} ms; per tile of CTA 0 (us since first stamp): "
          "(MMA0 warp | MMA1 warp || epilogue group)")
    for i in list(range(0, 12)) + list(range(100, 112)):
        r = [(int(x) - t0) / 1.9e3 if int(x) else float("nan") for x in t[i]]   # SM clocks at ~1.9 GHz -> us
        print(f"tile {i:3d} g{i % 4}: region free {r[0]:8.2f} | stage full {r[1]:8.2f} | MMA0 issued {r[2]:8.2f} | hid_ready seen "
              f"{r[5]:8.2f} || epilogue: acc0 seen {r[6]:8.2f} | acc1 seen {r[3]:8.2f} | LayerNorm done {r[4]:8.2f} | sums written {r[11]:8.2f} | tile done {r[7]:8.2f} || producer 0: e0 slots free {r[8]:8.2f} "
              f"gather slot free {r[9]:8.2f} issued {r[10]:8.2f} | producer 1: free {r[13]:8.2f} issued {r[14]:8.2f}")
L.check(L.lib().sgnn_debug_ws_flags(0))
