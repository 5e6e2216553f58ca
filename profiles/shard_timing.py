"""Sharded 1M-particle rollout under torchrun: ms per step for several rollout lengths and exchange modes.
torchrun --nproc-per-node N profiles/shard_timing.py"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

if __name__ == "__main__":
    from sgnn_b200 import sharding as SH
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=dev)
    traj, stats, radius = bench.make_case(seed=7, nx=2048, ny=512, nsteps=1)
    n = 2048 * 512
    lo, hi = SH.slab_bounds(n, world)[rank]
    init = traj["positions"][lo:hi, :bench.S].to(dev).contiguous()
    sim = bench.make_sim(stats, radius, dev)
    comm = SH.DistComm()
    SH.sharded_rollout(sim, init, 3, comm, n_layers=bench.LAYERS)
    for mode in ("graph", "eager_exchange", "host_plan"):
        os.environ.pop("SGNN_SHARD_EAGER_EXCHANGE", None)
        kw = {}
        if mode == "eager_exchange":
            os.environ["SGNN_SHARD_EAGER_EXCHANGE"] = "1"
        if mode == "host_plan":
            kw["fixed_ghosts"] = False
        for nsteps in (20, 50):
            dist.barrier()
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            SH.sharded_rollout(sim, init, nsteps, comm, n_layers=bench.LAYERS, refresh_every=10, **kw)
            b.record()
            torch.cuda.synchronize()
            ms = torch.tensor([a.elapsed_time(b)], device=dev, dtype=torch.float64)
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
            if rank == 0:
                print(f"{world} ranks {mode:15s} {nsteps:4d} steps: {ms.item() / nsteps:.3f} ms per step", flush=True)
    dist.destroy_process_group()
