"""Spatially sharded rollout of the single-scale simulator (BASELINE configs[4]: 1M+ particles).

The reference has no distributed code; this is new.  Design:

  * particles are partitioned into `world` contiguous GLOBAL-ID ranges ("slabs": the Taylor lattice
    is x-major, so id ranges are x-slabs).  Rank r owns ids [lo_r, hi_r).
  * a rank steps its owned particles plus a GHOST region: the particles of its two neighbours
    whose x lies within G = L * radius + skin of the slab's x-extent (L message-passing layers ->
    L hops of receptive field; particles in the outermost ring only contribute their encoder
    output).  Everything inside the ghost region is recomputed redundantly, so a step needs ONE
    exchange -- the new frame of the ghost particles, a few KB per neighbour over NVLink
    (torch.distributed P2P on NCCL) -- instead of one latent exchange per layer.
  * local arrays are ordered [left ghosts | owned | right ghosts], each part ascending in global id.
    Left ghosts have smaller ids than every owned particle, right ghosts larger, so LOCAL index
    order == GLOBAL id order and the reference's "first 20 sources in ascending index order"
    truncation (torch_cluster radius kernel) selects the same neighbours as the unsharded run:
    edge sets of owned particles are identical, bit for bit.
  * the ghost membership is refreshed every `refresh_every` steps (the `skin` absorbs the motion in
    between); between refreshes message sizes are fixed and no host synchronisation happens.

`Comm` abstracts the transport: `DistComm` (torch.distributed: NCCL on GPUs, gloo in the CPU tests)
and `ThreadComm` (W ranks emulated as threads of one process on ONE GPU, used by the single-GPU
parity test -- ranks that wait on each other must not be separate processes on one GPU).
"""
import threading
from typing import List, Optional, Tuple

import torch
import torch.distributed as dist


def slab_bounds(n: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous id ranges with equal counts (+-1)."""
    base, rem = divmod(n, world)
    out, lo = [], 0
    for r in range(world):
        hi = lo + base + (1 if r < rem else 0)
        out.append((lo, hi))
        lo = hi
    return out


def halo_width(n_layers: int, radius: float, skin: float) -> float:
    return n_layers * radius + skin


def select_for_neighbour(x_last_owned: torch.Tensor, nb_lo: float, nb_hi: float, g: float) -> torch.Tensor:
    """Ascending owned-local indices of the particles whose x lies within g of [nb_lo, nb_hi]."""
    m = (x_last_owned >= nb_lo - g) & (x_last_owned <= nb_hi + g)
    return torch.nonzero(m, as_tuple=False).flatten()


# ------------------------------------------------------------------------------------ transports
class Comm:
    rank: int
    world: int

    def allgather(self, t: torch.Tensor) -> List[torch.Tensor]:
        raise NotImplementedError

    def exchange(self, to_left: Optional[torch.Tensor], to_right: Optional[torch.Tensor],
                 from_left_like: Optional[torch.Tensor], from_right_like: Optional[torch.Tensor]):
        """Send to the two neighbours, receive into tensors shaped like the *_like arguments
        (None where there is no neighbour).  Returns (from_left, from_right)."""
        raise NotImplementedError


class DistComm(Comm):
    def __init__(self, group=None):
        self.group = group
        self.rank, self.world = dist.get_rank(group), dist.get_world_size(group)

    def allgather(self, t):
        out = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(out, t.contiguous(), group=self.group)
        return out

    def exchange(self, to_left, to_right, from_left_like, from_right_like):
        ops, fl, fr = [], None, None
        if self.rank > 0:
            fl = torch.empty_like(from_left_like)
            if to_left.numel():
                ops.append(dist.P2POp(dist.isend, to_left.contiguous(), self.rank - 1, self.group))
            if fl.numel():
                ops.append(dist.P2POp(dist.irecv, fl, self.rank - 1, self.group))
        if self.rank < self.world - 1:
            fr = torch.empty_like(from_right_like)
            if to_right.numel():
                ops.append(dist.P2POp(dist.isend, to_right.contiguous(), self.rank + 1, self.group))
            if fr.numel():
                ops.append(dist.P2POp(dist.irecv, fr, self.rank + 1, self.group))
        if ops:
            for req in dist.batch_isend_irecv(ops):
                req.wait()
        return fl, fr


class ThreadComm(Comm):
    """W ranks as threads of one process (single-GPU emulation / tests)."""

    class Shared:
        def __init__(self, world):
            self.world = world
            self.barrier = threading.Barrier(world)
            self.slots = [dict() for _ in range(world)]

    def __init__(self, shared: "ThreadComm.Shared", rank: int):
        self.sh, self.rank, self.world = shared, rank, shared.world

    def allgather(self, t):
        self.sh.slots[self.rank]["ag"] = t.clone()
        self.sh.barrier.wait()
        out = [self.sh.slots[r]["ag"] for r in range(self.world)]
        self.sh.barrier.wait()
        return out

    def exchange(self, to_left, to_right, from_left_like, from_right_like):
        if torch.cuda.is_available():
            torch.cuda.synchronize()
        self.sh.slots[self.rank]["l"] = None if to_left is None else to_left.clone()
        self.sh.slots[self.rank]["r"] = None if to_right is None else to_right.clone()
        self.sh.barrier.wait()
        fl = self.sh.slots[self.rank - 1]["r"] if self.rank > 0 else None
        fr = self.sh.slots[self.rank + 1]["l"] if self.rank < self.world - 1 else None
        self.sh.barrier.wait()
        return fl, fr


# ------------------------------------------------------------------------------------ ghost plan
class GhostPlan:
    """Who is a ghost of whom, fixed until the next refresh."""

    def __init__(self, comm: Comm, x_last_owned: torch.Tensor, g: float):
        dev = x_last_owned.device
        ext = torch.stack([x_last_owned.min(), x_last_owned.max()]) if x_last_owned.numel() else \
            torch.tensor([float("inf"), float("-inf")], device=dev)
        exts = comm.allgather(ext)
        r, w = comm.rank, comm.world
        # Every rank holds the same `exts`, so every rank takes the same decision here (a ValueError on one rank only
        # would leave the others hanging in the next collective).  Ghosts are exchanged with the two id-neighbours
        # only: slabs two or more apart must be farther than g apart in x, and an inner slab must be wider than g.
        lo_hi = [(float(e[0]), float(e[1])) for e in exts]
        for a in range(w):
            if 0 < a < w - 1 and lo_hi[a][1] - lo_hi[a][0] < g:
                raise ValueError(f"slab of rank {a} is narrower ({lo_hi[a][1] - lo_hi[a][0]:.3g}) than the halo "
                                 f"({g:.3g}): use fewer ranks")
            for b in range(a + 2, w):
                if not (lo_hi[b][0] - lo_hi[a][1] > g or lo_hi[a][0] - lo_hi[b][1] > g):
                    raise ValueError(f"slabs of ranks {a} and {b} are closer than the halo ({g:.3g}) in x: particle ids "
                                     "are not x-major (or the system deformed too far); re-partition before sharding")
        empty = torch.empty(0, dtype=torch.long, device=dev)
        self.send_left = select_for_neighbour(x_last_owned, float(exts[r - 1][0]), float(exts[r - 1][1]), g) \
            if r > 0 else empty
        self.send_right = select_for_neighbour(x_last_owned, float(exts[r + 1][0]), float(exts[r + 1][1]), g) \
            if r < w - 1 else empty
        counts = comm.allgather(torch.tensor([self.send_left.numel(), self.send_right.numel()],
                                             dtype=torch.long, device=dev))
        self.n_left = int(counts[r - 1][1]) if r > 0 else 0         # what my left neighbour sends right
        self.n_right = int(counts[r + 1][0]) if r < w - 1 else 0
        self.n_owned = x_last_owned.numel()

    @property
    def n_local(self) -> int:
        return self.n_left + self.n_owned + self.n_right


def exchange_rows(comm: Comm, plan: GhostPlan, owned_rows: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """owned_rows [n_owned, ...] -> (rows of my left ghosts, rows of my right ghosts)."""
    tail = owned_rows.shape[1:]
    like = lambda k: owned_rows.new_empty((k,) + tuple(tail))  # noqa: E731
    fl, fr = comm.exchange(owned_rows[plan.send_left] if comm.rank > 0 else None,
                           owned_rows[plan.send_right] if comm.rank < comm.world - 1 else None,
                           like(plan.n_left), like(plan.n_right))
    if fl is None:
        fl = like(0)
    if fr is None:
        fr = like(0)
    return fl, fr


# ------------------------------------------------------------------------------------ rollout
@torch.no_grad()
def sharded_rollout(simulator, positions_owned: torch.Tensor, nsteps: int, comm: Comm,
                    n_layers: int, skin: Optional[float] = None, refresh_every: int = 10,
                    return_edges: bool = False, use_graph: bool = True,
                    particle_types_owned: Optional[torch.Tensor] = None):
    """Autoregressive rollout of this rank's slab.

    positions_owned [n_owned, S, dim]: the initial window of the owned particles (global ids
    lo..hi-1 in order).  Returns (pred [nsteps, n_owned, dim], strain [nsteps, n_owned]) and, when
    return_edges, the step-0 edge list of the owned receivers as GLOBAL-id-relative local pairs
    (src_local - n_left + lo is the global id; see tests).

    Between two ghost refreshes the local system has a fixed size, so its step (neighbour search +
    network, 21 launches) runs on ONE window buffer [n_local, S + 1, dim] with fixed addresses -- frames
    [0, S) in, frame S out, then the window is shifted by a small device copy -- and is captured into a
    CUDA graph once per refresh and replayed every step (use_graph).  Launching the 21 kernels from
    Python costs ~0.5 ms per step, as much as the step itself on an eighth of a 1M-particle system
    (8 GPUs, r01n: 1.41 ms per step launched eagerly)."""
    dev = positions_owned.device
    no, s, dim = positions_owned.shape
    radius = float(simulator._connectivity_radius)
    skin = radius * 0.5 if skin is None else skin
    g = halo_width(n_layers, radius, skin)
    window = positions_owned.float().contiguous().clone()              # [no, S, dim], slides
    pred = torch.empty(nsteps, no, dim, dtype=torch.float32, device=dev)
    strain = torch.empty(nsteps, no, dtype=torch.float32, device=dev)
    edges0 = None
    k = 0
    while k < nsteps:
        m = min(refresh_every, nsteps - k)
        plan = GhostPlan(comm, window[:, -1, 0], g)
        gl, gr = exchange_rows(comm, plan, window)                     # full windows of the ghosts
        nl, nr, nloc = plan.n_left, plan.n_right, plan.n_local
        win = torch.empty(nloc, s + 1, dim, dtype=torch.float32, device=dev)
        win[:nl, :s], win[nl:nl + no, :s], win[nl + no:, :s] = gl, window, gr
        shift = torch.empty(nloc, s, dim, dtype=torch.float32, device=dev)
        types = torch.zeros(nloc, dtype=torch.int64, device=dev)
        if particle_types_owned is not None:                          # ghosts carry their owners' types
            tl, tr = exchange_rows(comm, plan, particle_types_owned.to(dev).long().contiguous())
            types[:nl], types[nl:nl + no], types[nl + no:] = tl, particle_types_owned.to(dev).long(), tr
        ptr = torch.tensor([0, nloc], dtype=torch.int32, device=dev)
        st_buf = torch.empty(nloc, dtype=torch.float32, device=dev)

        if k == 0:       # kernel plans and latent buffers with headroom: the local system changes size at every refresh
            simulator._reserve(nloc + max(nloc // 16, 1024), dev)

        def step():
            return simulator._step(win[:, :s], None, types, next_pos=win[:, s], strain=st_buf, example_ptr=ptr)

        graph = None
        for j in range(m):
            if use_graph and graph is None and m - j > 1 and (j == 1 or (j == 0 and k > 0)):
                # the very first step of the rollout runs eagerly (lazy initialisation, the step-0 edge list);
                # every other step replays the step captured once per refresh
                simulator._prepare(nloc, dev)                          # never build kernel plans inside a capture
                graph = torch.cuda.CUDAGraph()
                cap = torch.cuda.Stream(device=dev)                    # (torch.cuda.graph() would also empty the
                cap.wait_stream(torch.cuda.current_stream())           #  allocator cache at every refresh)
                with torch.cuda.stream(cap):
                    graph.capture_begin(capture_error_mode="thread_local")
                    held = step()      # noqa: F841 -- the engine / edge set whose pointers the graph replays stay alive with it
                    graph.capture_end()
                torch.cuda.current_stream().wait_stream(cap)
            if graph is not None:
                graph.replay()
            else:
                _, csr = step()
                if return_edges and edges0 is None:
                    ei = csr.edge_index()
                    keep = (ei[1] >= nl) & (ei[1] < nl + no)
                    edges0 = (ei[:, keep], nl)
            new_owned = win[nl:nl + no, s]
            pred[k + j] = new_owned
            strain[k + j] = st_buf[nl:nl + no]
            if j + 1 < m:                                              # ghosts take their owners' values
                fl, fr = exchange_rows(comm, plan, new_owned.contiguous())
                win[:nl, s] = fl
                win[nl + no:, s] = fr
            shift.copy_(win[:, 1:])                                    # slide the window: frames 1..S -> 0..S-1
            win[:, :s] = shift
        window = win[nl:nl + no, :s].contiguous()
        k += m
    if return_edges:
        return pred, strain, edges0
    return pred, strain
