"""In-kernel timeline of the generic node_kernel (CTA 0) in a multi-scale step (BASELINE configs[3], 4800 particles,
H = 128, two hidden layers, streamed weights).  Needs a SGNN_TIMELINE=1 build of the library."""
import ctypes as C
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from sgnn_b200 import _lib as L  # noqa: E402
from sgnn_b200 import synthetic  # noqa: E402
from sgnn_b200.multi_scale import MultiScaleConfig, MultiScaleGraph, MultiScaleSimulator  # noqa: E402
from sgnn_b200.single_scale import evaluate as EV  # noqa: E402

dev = torch.device("cuda", 0)
S = bench.S
traj = synthetic.taylor_trajectory(bench.NX, bench.NY, 0.5, nframes=S + 4, seed=0)
stats = synthetic.normalization_stats(synthetic.metadata_from(traj["positions"]), 0.02)
torch.manual_seed(1)
sim = MultiScaleSimulator(2, (S - 1) * 2 + 1, 3, 128, 128, bench.LAYERS, 2, stats, 1, 9, num_scales=2,
                          window_size=2, radius_multiplier=2.0, device=dev).to(dev).eval()
pos = traj["positions"].to(dev)
sim.set_static_graph(MultiScaleGraph(MultiScaleConfig(2, 2, 2.0)).create_all_edges(pos[:, 0].contiguous()))
n, t, dim = pos.shape
plan = EV.RolloutPlan(sim, n, t, S, dim, "autoregressive", dev, use_graph=False)
plan.traj[:, :S].copy_(pos[:, :S])
plan.launch_steps(0, 2)
torch.cuda.synchronize()
buf = torch.zeros(32, dtype=torch.int64, device=dev)
ebuf = torch.zeros(16, dtype=torch.int64, device=dev)
L.check(L.lib().sgnn_debug_node_generic_timeline(C.c_void_p(buf.data_ptr())), 0)
L.check(L.lib().sgnn_debug_edge_generic_timeline(C.c_void_p(ebuf.data_ptr())), 0)
plan.launch_steps(2, 3)
torch.cuda.synchronize()
L.check(L.lib().sgnn_debug_node_generic_timeline(None), 0)
L.check(L.lib().sgnn_debug_edge_generic_timeline(None), 0)
e = ebuf.cpu().tolist()
enames = ["entry", "setup + tmem alloc + first prefetch", "pdl wait + edge count + vectors", "e0 tile + GEMM 0", "P[tgt] + Q[src] + ReLU",
          "hidden GEMMs and epilogues", "output GEMM", "LayerNorm -> message tile", "segmented sums", "stitch + teardown"]
print("last edge_layer_kernel of the step")
for k in range(1, 10):
    print(f"  {enames[k]:40s} +{(e[k] - e[k - 1]) / 1e3:7.2f} us")
print(f"  total {(e[9] - e[0]) / 1e3:.2f} us")
if e[10]:
    print(f"  inside the sums (thread 0): messages in registers +{(e[10] - e[7]) / 1e3:.2f}, walk + stores +{(e[11] - e[10]) / 1e3:.2f}, "
          f"barrier +{(e[8] - e[11]) / 1e3:.2f} us")
t = buf.cpu().tolist()
names = ["entry", "setup + tmem alloc + first prefetch", "pdl wait + vectors landed", "prologue (agg | x rows)",
         "GEMM 0 (K = 256)", "epilogue 0", "hidden + output GEMMs and epilogues", "LayerNorm moments", "affine + residual + store",
         "tail GEMMs", "tail epilogue / stores", "teardown"]
for base, label in ((0, "last PQ-tail kernel of the step"), (16, "DECODE-tail kernel")):
    print(label)
    for k in range(1, 12):
        if t[base + k] and t[base + k - 1]:
            print(f"  {names[k]:40s} +{(t[base + k] - t[base + k - 1]) / 1e3:7.2f} us")
    print(f"  total {(t[base + 11] - t[base]) / 1e3:.2f} us")
