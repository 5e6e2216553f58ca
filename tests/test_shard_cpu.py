"""CPU (gloo, world_size 2 and 3): ghost selection + halo exchange of sgnn_b200/sharding.py.
The local particle set [left ghosts | owned | right ghosts] of every rank must reproduce, for each
OWNED receiver, exactly the neighbour list of the unsharded search (oracle radius graph, first-K
truncation in ascending global index order), and per-step exchanges must deliver the owners' rows."""
import os

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _positions(nx=40, ny=9, spacing=5.0, seed=0):
    from sgnn_b200 import synthetic
    p = synthetic.taylor_bar_lattice(nx, ny, spacing)
    g = torch.Generator().manual_seed(seed)
    return p + 0.05 * spacing * torch.randn(p.shape, generator=g)


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle.radius import radius_graph_np
    from sgnn_b200 import sharding as S
    pos = _positions()
    n = pos.shape[0]
    radius, layers, K = 15.0, 1, 20
    lo, hi = S.slab_bounds(n, world)[rank]
    comm = S.DistComm()
    plan = S.GhostPlan(comm, pos[lo:hi, 0].contiguous(), S.halo_width(layers, radius, 2.0))
    ids = torch.arange(lo, hi).float()[:, None]
    payload = torch.cat([pos[lo:hi], ids], 1).contiguous()          # x, y, global id
    gl, gr = S.exchange_rows(comm, plan, payload)
    local = torch.cat([gl, payload, gr])
    gid = local[:, 2].long()
    assert torch.equal(gid, torch.sort(gid).values), "local order must be global-id order"
    ei = torch.from_numpy(radius_graph_np(local[:, :2].numpy().astype(np.float32), radius, None, True, K))
    keep = (ei[1] >= plan.n_left) & (ei[1] < plan.n_left + plan.n_owned)
    out[rank] = torch.stack([gid[ei[0][keep]], gid[ei[1][keep]]])
    # a second exchange with the same plan (per-step message): fixed sizes, owners' values
    fl, fr = S.exchange_rows(comm, plan, (payload * 2).contiguous())
    assert torch.equal(fl, gl * 2) and torch.equal(fr, gr * 2)
    dist.destroy_process_group()


def _run(world, port):
    from oracle.radius import radius_graph_np
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
    pos = _positions()
    full = torch.from_numpy(radius_graph_np(pos.numpy().astype(np.float32), 15.0, None, True, 20))
    got = torch.cat([out[r] for r in range(world)], 1)
    assert torch.equal(got, full), "sharded edge sets differ from the unsharded search"


def test_ghost_plan_reproduces_global_edges_world2():
    _run(2, 29621)


def test_ghost_plan_reproduces_global_edges_world3():
    _run(3, 29622)


def test_slab_bounds():
    from sgnn_b200 import sharding as S
    b = S.slab_bounds(10, 3)
    assert b == [(0, 4), (4, 7), (7, 10)]
