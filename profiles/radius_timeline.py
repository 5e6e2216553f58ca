"""Phase stamps of radius_cluster_kernel (CTA 0, thread 0) at N=4800 under CUDA-graph replay (SGNN_TIMELINE build)."""
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
steps = 20
traj, stats, radius = bench.make_case(seed=0, nsteps=steps)
sim = bench.make_sim(stats, radius, dev)
n = 4800
plan = EV.RolloutPlan(sim, n, bench.S + steps, bench.S, 2, "one_step", dev, use_graph=False)
plan.traj.copy_(traj["positions"].to(dev))
plan.launch_steps(0, 2)
torch.cuda.synchronize()
buf = torch.zeros(16, dtype=torch.int64, device=dev)
L.check(L.lib().sgnn_debug_radius_timeline(C.c_void_p(buf.data_ptr())), 0)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    plan.launch_steps(0, steps)
for _ in range(3):
    g.replay()
torch.cuda.synchronize()
L.check(L.lib().sgnn_debug_radius_timeline(None), 0)
t = buf.cpu().tolist()
names = ["entry", "pdl_wait returned", "A box partials delivered", "  cluster barrier", "B cells counted", "  cluster barrier",
         "C block scan + totals", "  cluster barrier", "C offsets applied", "  cluster barrier (after D fill below)",
         "E search done", "  cluster barrier", "F row pointers + edges written"]
names[8], names[9] = "C offsets applied + barrier + D fill", "  cluster barrier"
for k in range(13):
    print(f"{names[k]:42s} {(t[k] - t[0]) / 1965.0:8.2f} us")
