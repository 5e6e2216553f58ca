"""Short un-graphed run of the single-scale step for ncu (launch list / --set full captures).

    python profiles/profile_step.py --nx 2048 --ny 512 --steps 2

Runs `steps` rollout steps (neighbour search + encoders + L layers + decoder/integrator) of
the bench model on a synthetic Taylor bar, each kernel a separate launch (no CUDA graph), and
prints the CUDA-event time of the last step.  Not a benchmark: bench.py is.
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nx", type=int, default=2048)
    ap.add_argument("--ny", type=int, default=512)
    ap.add_argument("--steps", type=int, default=2)
    ap.add_argument("--regime", default="A")
    args = ap.parse_args()
    from sgnn_b200.single_scale import evaluate as EV
    bench.REGIME = args.regime
    dev = torch.device("cuda", 0)
    traj, stats, radius = bench.make_case(seed=7, nx=args.nx, ny=args.ny, nsteps=args.steps)
    sim = bench.make_sim(stats, radius, dev)
    n = args.nx * args.ny
    S = bench.S
    plan = EV.RolloutPlan(sim, n, S + args.steps, S, 2, "autoregressive", dev, use_graph=False)
    plan.traj[:, :S].copy_(traj["positions"][:, :S].to(dev))
    torch.cuda.synchronize()
    for k in range(args.steps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        plan.launch_steps(k, k + 1)
        b.record()
        torch.cuda.synchronize()
        print(f"step {k}: {a.elapsed_time(b):.3f} ms")
    _, rg = sim._engine(n, dev)
    print("N", n, "E", int(rg.csr.rowptr[rg.csr.n_nodes].item()))


if __name__ == "__main__":
    main()
