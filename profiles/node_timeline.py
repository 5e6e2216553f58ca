"""In-kernel timeline of node_fast_kernel at N=4800 (one tile per CTA): where do ~15 us go?"""
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
traj, stats, radius = bench.make_case(seed=0, nsteps=4)
sim = bench.make_sim(stats, radius, dev)
n = 4800
plan = EV.RolloutPlan(sim, n, bench.S + 4, bench.S, 2, "autoregressive", dev, use_graph=False)
plan.traj[:, :bench.S].copy_(traj["positions"][:, :bench.S].to(dev))
plan.launch_steps(0, 2)
torch.cuda.synchronize()
buf = torch.zeros(16, dtype=torch.int64, device=dev)
L.check(L.lib().sgnn_debug_node_timeline(C.c_void_p(buf.data_ptr())), 0)
plan.launch_steps(2, 3)
torch.cuda.synchronize()
t = buf.cpu().tolist()
names = ["entry", "setup+tmem alloc", "weights landed", "pdl wait", "agg row loaded", "mma agg", "x row loaded+st",
         "mma x", "mlp layers", "ln+residual+store", "tail", "dealloc/exit"]
for k in range(1, 12):
    print(f"{names[k]:22s} +{(t[k] - t[k-1]) / 1e3:7.2f} us")
print("total", (t[11] - t[0]) / 1e3, "us (last node kernel of the step: DECODE tail)")
L.check(L.lib().sgnn_debug_node_timeline(None), 0)
