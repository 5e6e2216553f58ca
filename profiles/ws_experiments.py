"""A/B timings of the gather-mode edge layer at ~1M particles (measurement switches of
edge_layer_ws.cu; results of the flagged runs are wrong by construction)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
from sgnn_b200 import _lib as L  # noqa: E402


def main():
    from sgnn_b200.single_scale import evaluate as EV
    dev = torch.device("cuda", 0)
    nx, ny = 2048, 512
    traj, stats, radius = bench.make_case(seed=7, nx=nx, ny=ny, nsteps=2)
    sim = bench.make_sim(stats, radius, dev)
    n = nx * ny
    plan = EV.RolloutPlan(sim, n, bench.S + 2, bench.S, 2, "autoregressive", dev, use_graph=False)
    plan.traj[:, :bench.S].copy_(traj["positions"][:, :bench.S].to(dev))
    plan.launch_steps(0, 1)
    torch.cuda.synchronize()
    for flags in [int(a) for a in (sys.argv[1:] or ["0", "1"])]:
        L.check(L.lib().sgnn_debug_ws_flags(flags))
        ms, e = bench.measure_edge_layer(sim, n, dev, reps=20)
        print(f"flags {flags}: edge layer {ms:.4f} ms  (E={e})")
    L.check(L.lib().sgnn_debug_ws_flags(0))


if __name__ == "__main__":
    main()
