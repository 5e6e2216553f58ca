"""In-kernel timeline of node_ws_kernel (CTA 0, last PX layer) at N=4800 under CUDA-graph replay: what
the ~9.7 us per node-layer kernel consist of.  Needs an SGNN_TIMELINE=1 build:
    SGNN_TIMELINE=1 python -c "from sgnn_b200 import build_lib; build_lib.build(force=True)"
"""
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
buf = torch.zeros(32, dtype=torch.int64, device=dev)
L.check(L.lib().sgnn_debug_node_ws_timeline(C.c_void_p(buf.data_ptr())), 0)
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    plan.launch_steps(0, steps)
for _ in range(3):
    g.replay()
torch.cuda.synchronize()
t = buf.cpu().tolist()
names = ["entry", "prologue done (barriers, TMEM, weight TMA issued)", "pdl_wait returned", "tile [agg|x] landed",
         "acc0 ready (patch + GEMM 0)", "acc1 ready (hidden epilogue + GEMM 1)", "LN + residual done, x' to TMEM",
         "acc2 ready (tail GEMM P)", "P stored, TMA stores drained", "all warps done",
         "  epilogue: agg row patched", "  MMA0 warp: patched seen", "  MMA0 warp: MMAs + commit issued",
         "  MMA0 warp: weights landed"]
print("slot                                                 globaltimer us   clock64 us (1.965 GHz)")
for k in range(14):
    print(f"{names[k]:52s} {(t[k] - t[0]) / 1e3:10.2f} {(t[16 + k] - t[16]) / 1965.0:14.2f}")
L.check(L.lib().sgnn_debug_node_ws_timeline(None), 0)
