"""A few un-captured steps of the training engine on BASELINE configs[2] (for ncu: launch lists / full captures).
python profiles/train_engine_step.py [fp32|tf32] [steps]"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

if __name__ == "__main__":
    from sgnn_b200 import train_engine, training
    prec = sys.argv[1] if len(sys.argv) > 1 else "fp32"
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    dev = torch.device("cuda", 0)
    batch, stats, radius = bench.make_train_batch(dev, 8, seed0=100)
    sim = bench.make_sim(stats, radius, dev).train()
    eng = train_engine.TrainEngine(sim, training.make_optimizer(sim.parameters(), 1e-4), batch["positions"].shape[0], dev,
                                   precision=prec, use_graph=False)
    gen = torch.Generator(device=dev).manual_seed(2)
    for _ in range(steps):
        eng.step(batch, 0.5, generator=gen)
    torch.cuda.synchronize()
    print("ok")
