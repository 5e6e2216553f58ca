"""Kernel start / duration inside the CUDA-graph replay of a multi-scale rollout (BASELINE configs[3]) from CUPTI
(torch.profiler): where a 0.58 ms step goes -- kernel time or the gaps between kernels.
python profiles/ms_graph_trace.py [steps]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from sgnn_b200 import synthetic  # noqa: E402
from sgnn_b200.multi_scale import MultiScaleConfig, MultiScaleGraph, MultiScaleSimulator  # noqa: E402
from sgnn_b200.single_scale import evaluate as EV  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 6
dev = torch.device("cuda", 0)
S = bench.S
traj = synthetic.taylor_trajectory(bench.NX, bench.NY, 0.5, nframes=S + steps, seed=0)
stats = synthetic.normalization_stats(synthetic.metadata_from(traj["positions"]), 0.02)
torch.manual_seed(1)
sim = MultiScaleSimulator(2, (S - 1) * 2 + 1, 3, 128, 128, bench.LAYERS, 2, stats, 1, 9, num_scales=2,
                          window_size=2, radius_multiplier=2.0, device=dev).to(dev).eval()
pos = traj["positions"].to(dev)
sim.set_static_graph(MultiScaleGraph(MultiScaleConfig(2, 2, 2.0)).create_all_edges(pos[:, 0].contiguous()))
n, t, dim = pos.shape
EV.VERBOSE = False
types = torch.zeros(n, dtype=torch.int64, device=dev)
strains = traj["strains"].to(dev)
kw = dict(use_cuda_graph=True)
for _ in range(2):
    EV.rollout(sim, pos, types, torch.tensor(n), strains, steps, dim, dev, S, "autoregressive", **kw)
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    EV.rollout(sim, pos, types, torch.tensor(n), strains, steps, dim, dev, S, "autoregressive", **kw)
    torch.cuda.synchronize()
ev = sorted([e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA and "sgnn" in e.name],
            key=lambda e: e.time_range.start)
per = len(ev) // steps
print(f"{len(ev)} sgnn kernels, {per} per step")
seg = ev[per * (steps - 2):per * (steps - 1)]          # one step in the middle of the replay
t0 = seg[0].time_range.start
busy = 0.0
prev_end = t0
for e in seg:
    st, en = e.time_range.start, e.time_range.end
    print(f"{e.name[:48]:50s} start {st - t0:8.1f} us  dur {en - st:6.1f}  gap {st - prev_end:6.1f}")
    busy += en - st
    prev_end = max(prev_end, en)
print(f"step span {prev_end - t0:.1f} us, kernel time {busy:.1f} us")
