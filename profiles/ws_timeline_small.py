"""In-kernel timeline of edge_layer_ws_kernel (CTA 0) at N=4800 under CUDA-graph replay (SGNN_TIMELINE build):
what the ~9.2 us per edge-layer kernel consist of.  SM-clock stamps, us at 1.965 GHz since kernel entry."""
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
buf = torch.zeros(256 * 16, dtype=torch.int64, device=dev)
L.check(L.lib().sgnn_debug_ws_timeline(C.c_void_p(buf.data_ptr())), 0)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    plan.launch_steps(0, steps)
for _ in range(3):
    g.replay()
torch.cuda.synchronize()
L.check(L.lib().sgnn_debug_ws_timeline(None), 0)
t = buf.cpu().view(256, 16)
t0 = int(t[255, 0])
us = lambda x: (int(x) - t0) / 1965.0 if int(x) else float("nan")  # noqa: E731
k = t[255]
print(f"kernel: entry 0.00 | prologue done {us(k[1]):.2f} | pdl_wait returned {us(k[2]):.2f} | n_edges loaded {us(k[3]):.2f} | "
      f"all warps done {us(k[4]):.2f}")
for i in range(2):
    r = [us(x) for x in t[i]]
    print(f"tile {i}: producer0 slots free {r[8]:.2f} gather slot free {r[9]:.2f} issued {r[10]:.2f} | producer1 free {r[13]:.2f} "
          f"issued {r[14]:.2f} | MMA0: region free {r[0]:.2f} stage full {r[1]:.2f} issued {r[2]:.2f} | MMA1: hid_ready seen {r[5]:.2f} "
          f"| epilogue: acc0 seen {r[6]:.2f} acc1 seen {r[3]:.2f} LN done {r[4]:.2f} warp 0 sums done {r[11]:.2f} group done {r[7]:.2f}")
