"""CPU (gloo, world_size 2): the data-parallel plumbing of sgnn_b200/training.py -- flat gradient
SUM-allreduce with a parameter that gets no gradient, global particle count, LR schedule."""
import os

import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from sgnn_b200 import training
    torch.manual_seed(0)
    lin = torch.nn.Linear(4, 3)
    unused = torch.nn.Embedding(2, 5)                      # never receives a gradient
    params = list(lin.parameters()) + list(unused.parameters())
    x = torch.arange(8, dtype=torch.float32).reshape(2, 4) + rank
    n_glob = training.global_particle_count(2, torch.device("cpu"))
    (lin(x).square().sum() / n_glob).backward()
    training.FlatGradAllReduce(params)()
    out[rank] = (n_glob, lin.weight.grad.clone(), unused.weight.grad.clone())
    dist.destroy_process_group()


def test_flat_grad_allreduce_matches_single_process():
    world = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, 29611, out), nprocs=world, join=True)
    torch.manual_seed(0)
    lin = torch.nn.Linear(4, 3)
    x = torch.cat([torch.arange(8, dtype=torch.float32).reshape(2, 4) + r for r in range(world)])
    (lin(x).square().sum() / 4).backward()
    for r in range(world):
        n_glob, g, gu = out[r]
        assert n_glob == 4
        assert torch.allclose(g, lin.weight.grad, atol=1e-5)
        assert float(gu.abs().max()) == 0.0


def test_learning_rate_schedule():
    from sgnn_b200 import training
    assert abs(training.learning_rate(0, 1e-4, 0.1, 5e6) - (1e-4 + 1e-6)) < 1e-12
    assert abs(training.learning_rate(5_000_000, 1e-4, 0.1, 5e6) - (1e-5 + 1e-6)) < 1e-12
