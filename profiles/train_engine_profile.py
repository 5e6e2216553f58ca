"""Training step (sgnn_b200/train_engine.py) on BASELINE configs[2] (8 x 4800 particles): steps/s of the captured
step and the per-family CUDA-event breakdown of one un-captured step.  python profiles/train_engine_profile.py [fp32|tf32]"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

if __name__ == "__main__":
    prec = sys.argv[1] if len(sys.argv) > 1 else "fp32"
    dev = torch.device("cuda", 0)
    out = bench._train_config(dev, 0, 1, 30, 5, 8, prec)
    bd = out.pop("breakdown_ms")
    print(json.dumps(out))
    tot = bd.get("forward_backward_total", {}).get("total_ms", 0.0)
    for k, v in sorted(bd.items(), key=lambda kv: -kv[1]["total_ms"]):
        print(f"{k:26s} {v['launch_groups']:4d} groups {v['total_ms']:9.3f} ms {100 * v['total_ms'] / max(tot, 1e-9):6.1f} %")
    # kernel durations of un-captured steps (CUPTI through torch.profiler): the per-family events above include
    # the host time between two launches
    from torch.profiler import ProfilerActivity, profile
    from sgnn_b200 import train_engine, training
    batch, stats, radius = bench.make_train_batch(dev, 8, seed0=100)
    sim = bench.make_sim(stats, radius, dev).train()
    eng = train_engine.TrainEngine(sim, training.make_optimizer(sim.parameters(), 1e-4), batch["positions"].shape[0], dev,
                                   precision=prec, use_graph=False)
    gen = torch.Generator(device=dev).manual_seed(2)
    for _ in range(2):
        eng.step(batch, 0.5, generator=gen)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        for _ in range(3):
            eng.step(batch, 0.5, generator=gen)
        torch.cuda.synchronize()
    ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    print(f"GPU kernel time {sum(e.device_time for e in ev if e.device_time) / 3e3:.3f} ms per step over {len(ev) / 3:.0f} kernels")
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=30, max_name_column_width=60))
