"""torch.profiler table of the data-parallel train step (BASELINE configs[2], one GPU):
    python profiles/train_profile.py
Prints wall ms per step, the sum of GPU kernel time per step and the top kernels."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402


def main():
    from sgnn_b200 import training
    mm = sys.argv[1] if len(sys.argv) > 1 else "fp32"
    dev = torch.device("cuda", 0)
    batch, stats, radius = bench.make_train_batch(dev, 8, seed0=100)
    sim = bench.make_sim(stats, radius, dev).train()
    opt = torch.optim.Adam(sim.parameters(), lr=1e-4, fused=True)
    gen = torch.Generator(device=dev).manual_seed(2)
    n = batch["positions"].shape[0]
    for _ in range(3):
        training.train_step(sim, opt, batch, 0.5, n_global=n, generator=gen, matmul=mm)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        training.train_step(sim, opt, batch, 0.5, n_global=n, generator=gen, matmul=mm)
    b.record()
    torch.cuda.synchronize()
    print(mm, f"wall {a.elapsed_time(b) / 10:.3f} ms per step")
    from torch.profiler import profile, ProfilerActivity
    with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA]) as prof:
        for _ in range(3):
            training.train_step(sim, opt, batch, 0.5, n_global=n, generator=gen, matmul=mm)
        torch.cuda.synchronize()
    ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
    tot = sum(e.device_time for e in ev if e.device_time) / 3e3
    print(f"GPU kernel time {tot:.3f} ms per step over {len(ev) / 3:.0f} kernels")
    print(prof.key_averages().table(sort_by="cuda_time_total", row_limit=25, max_name_column_width=70))


if __name__ == "__main__":
    main()
