"""Un-graphed multi-scale rollout steps (BASELINE configs[3]) for ncu launch lists:
    ncu --metrics gpu__time_duration.sum ... python profiles/profile_multiscale.py --steps 2"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=2)
    args = ap.parse_args()
    from sgnn_b200 import synthetic
    from sgnn_b200.multi_scale import MultiScaleConfig, MultiScaleGraph, MultiScaleSimulator
    from sgnn_b200.single_scale import evaluate as EV
    dev = torch.device("cuda", 0)
    S = bench.S
    traj = synthetic.taylor_trajectory(bench.NX, bench.NY, 0.5, nframes=S + args.steps, seed=0)
    stats = synthetic.normalization_stats(synthetic.metadata_from(traj["positions"]), 0.02)
    torch.manual_seed(1)
    sim = MultiScaleSimulator(2, (S - 1) * 2 + 1, 3, 128, 128, bench.LAYERS, 2, stats, 1, 9, num_scales=2,
                              window_size=2, radius_multiplier=2.0, device=dev).to(dev).eval()
    pos = traj["positions"].to(dev)
    graph = MultiScaleGraph(MultiScaleConfig(2, 2, 2.0)).create_all_edges(pos[:, 0].contiguous())
    sim.set_static_graph(graph)
    n, t, dim = pos.shape
    plan = EV.RolloutPlan(sim, n, t, S, dim, "autoregressive", dev, use_graph=False)
    plan.traj[:, :S].copy_(pos[:, :S])
    torch.cuda.synchronize()
    for k in range(args.steps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        plan.launch_steps(k, k + 1)
        b.record()
        torch.cuda.synchronize()
        print(f"step {k}: {a.elapsed_time(b):.3f} ms")


if __name__ == "__main__":
    main()
