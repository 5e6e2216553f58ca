"""Where the 4800-particle rollout step spends its time under CUDA-graph replay.

    python profiles/chain_breakdown.py [--nx 120 --ny 40]

Captures 100 teacher-forced steps (one_step mode: the window always reads ground-truth frames, so
skipping kernels cannot degenerate the geometry) into a CUDA graph with only ONE kernel family
enabled (engine.SKIP holds the others) and times replays with CUDA events.  The per-family
microseconds do not add up to the full step exactly (launch overlap differs), they rank the chains.
"""
import argparse
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402

FAMILIES = ["radius_graph", "edge_encoder", "node_encoder", "edge_layer", "node_layer"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nx", type=int, default=120)
    ap.add_argument("--ny", type=int, default=40)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--mode", default="one_step", choices=["one_step", "autoregressive"],
                    help="autoregressive: full step only (skipping kernels would degenerate the geometry)")
    args = ap.parse_args()
    from sgnn_b200 import engine as ENG
    from sgnn_b200.single_scale import evaluate as EV
    dev = torch.device("cuda", 0)
    traj, stats, radius = bench.make_case(seed=7, nx=args.nx, ny=args.ny, nsteps=args.steps)
    sim = bench.make_sim(stats, radius, dev)
    n, S = args.nx * args.ny, bench.S
    plan = EV.RolloutPlan(sim, n, S + args.steps, S, 2, args.mode, dev, use_graph=False)
    plan.traj.copy_(traj["positions"].to(dev))
    plan.launch_steps(0, 2)
    torch.cuda.synchronize()
    flush = torch.empty(512 << 20, dtype=torch.uint8, device=dev)

    def timed(skip):
        ENG.SKIP = set(skip)
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            plan.launch_steps(0, args.steps)
        ENG.SKIP = set()
        best = 1e9
        for _ in range(args.reps):
            flush.fill_(1)
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            g.replay()
            b.record()
            torch.cuda.synchronize()
            best = min(best, a.elapsed_time(b))
        return best * 1e3 / args.steps

    print(f"N={n}  us per step under graph replay (best of {args.reps})")
    print(f"  full step            {timed([]):8.2f}")
    if args.mode == "autoregressive":
        _, rg = sim._engine(n, dev)
        print("  edges after the rollout", int(rg.csr.rowptr[rg.csr.n_nodes].item()))
        return
    for f in FAMILIES:
        print(f"  only {f:15s} {timed([x for x in FAMILIES if x != f]):8.2f}")
    for f in FAMILIES:
        print(f"  all but {f:12s} {timed([f]):8.2f}")


if __name__ == "__main__":
    main()
