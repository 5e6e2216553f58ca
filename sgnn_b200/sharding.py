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
import os
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


class FixedGhostPlan:
    """Ghost membership with FIXED message sizes: every rank sends each neighbour exactly `cap` particles -- the `cap`
    owned particles closest in x to that neighbour's extent, a superset of the particles within the halo width g as
    long as at most `cap` lie inside it.  Selection is a top-k + sort on the device, so a refresh neither reads the
    host nor changes the size of the local system: the CUDA graph captured for one refresh interval stays valid for the
    whole rollout.  refresh() reports how many particles the halos really needed, so the caller can re-plan when `cap` is too small."""

    def __init__(self, comm: Comm, n_owned: int, cap: int, g: float, device):
        r, w = comm.rank, comm.world
        self.comm, self.g, self.n_owned = comm, g, n_owned
        self.cap_l = cap if r > 0 else 0                    # what I send left = what my left neighbour receives
        self.cap_r = cap if r < w - 1 else 0
        self.n_left, self.n_right = self.cap_l, self.cap_r   # symmetric: every rank uses the same cap
        # persistent index buffers: a captured step reads them by address, a refresh rewrites them in place
        self.send_left = torch.zeros(self.cap_l, dtype=torch.long, device=device)
        self.send_right = torch.zeros(self.cap_r, dtype=torch.long, device=device)
        self.cap = cap

    @property
    def n_local(self) -> int:
        return self.n_left + self.n_owned + self.n_right

    def _closest(self, x, lo, hi, cap):
        d = torch.clamp(torch.maximum(lo - x, x - hi), min=0.0)               # distance in x to the neighbour's extent
        self._inside = torch.maximum(self._inside, (d <= self.g).sum())       # particles the halo really needs
        vals, idx = torch.topk(d, cap, largest=False, sorted=True)
        return torch.sort(idx).values                                         # ascending owned-local index = global-id order

    def refresh(self, x_last_owned: torch.Tensor) -> int:
        """Re-select the ghosts on the device.  Returns the largest number of particles any rank found inside a halo
        (ONE host read per refresh, after all device work is queued): the caller re-plans with a larger capacity when
        it exceeds `cap`."""
        ext = torch.stack([x_last_owned.min(), x_last_owned.max()])
        exts = self.comm.allgather(ext)                                       # device tensors
        self._inside = torch.zeros((), dtype=torch.long, device=x_last_owned.device)
        r = self.comm.rank
        if self.cap_l:
            self.send_left.copy_(self._closest(x_last_owned, exts[r - 1][0], exts[r - 1][1], self.cap_l))
        if self.cap_r:
            self.send_right.copy_(self._closest(x_last_owned, exts[r + 1][0], exts[r + 1][1], self.cap_r))
        return max(int(c) for c in self.comm.allgather(self._inside.reshape(1)))


def exchange_rows(comm: Comm, plan, owned_rows: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
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
                    particle_types_owned: Optional[torch.Tensor] = None, fixed_ghosts: bool = False,
                    ghost_headroom: float = 1.25):
    """Autoregressive rollout of this rank's slab.

    positions_owned [n_owned, S, dim]: the initial window of the owned particles (global ids
    lo..hi-1 in order).  Returns (pred [nsteps, n_owned, dim], strain [nsteps, n_owned]) and, when
    return_edges, the step-0 edge list of the owned receivers as GLOBAL-id-relative local pairs
    (src_local - n_left + lo is the global id; see tests).

    The local system [left ghosts | owned | right ghosts] steps on ONE window buffer [n_local, S + 1, dim] with
    fixed addresses -- frames [0, S) in, frame S out, then the window is shifted by a small device copy -- so its
    step (neighbour search + network, 21 launches) is captured into a CUDA graph and replayed (use_graph).

    Default: every refresh re-plans on the host (GhostPlan) and re-captures the step.  fixed_ghosts=True (opt-in): the
    FIRST refresh selects the ghosts on the host and fixes the message size at ghost_headroom x the largest count;
    every later refresh re-selects that many particles on the device (FixedGhostPlan: top-k by distance to the
    neighbour's extent, one host read of the needed count per refresh, re-planned when the halo outgrows the capacity).
    The local system then keeps its size, the step is captured once per rollout and -- over NCCL -- the ghost-frame
    exchange and the window shift are captured with it (SGNN_SHARD_EAGER_EXCHANGE=1 keeps them outside).  Measured on
    8 B200 at 1M particles (profiles/README.md r02e, ms per step for 20 / 50-step rollouts): host plan 1.29 / 1.18,
    fixed + eager exchange 1.14 / 1.67, fixed + captured exchange 1.41 / 2.09 -- the captured NCCL send / recv pairs
    cost more than the launches they replace and the top-k selection as much as the re-capture, so it is not the
    default."""
    dev = positions_owned.device
    no, s, dim = positions_owned.shape
    radius = float(simulator._connectivity_radius)
    skin = radius * 0.5 if skin is None else skin
    g = halo_width(n_layers, radius, skin)
    window = positions_owned.float().contiguous().clone()              # [no, S, dim], slides (host-planned mode)
    pred = torch.empty(nsteps, no, dim, dtype=torch.float32, device=dev)
    strain = torch.empty(nsteps, no, dtype=torch.float32, device=dev)
    edges0 = None
    fixed = None           # FixedGhostPlan once the message sizes are fixed
    win = graph = held = None
    k = 0
    while k < nsteps:
        m = min(refresh_every, nsteps - k)
        if fixed is not None:
            if fixed.refresh(win[nl:nl + no, s - 1, 0]) > fixed.cap:   # the halo outgrew the fixed capacity: re-plan on the host
                window, fixed = win[nl:nl + no, :s].contiguous(), None
        if fixed is None:
            plan = GhostPlan(comm, window[:, -1, 0], g)
            if fixed_ghosts and comm.world > 1:
                counts = comm.allgather(torch.tensor([plan.send_left.numel(), plan.send_right.numel(), no],
                                                     dtype=torch.long, device=dev))
                most = max(int(c[0]) for c in counts), max(int(c[1]) for c in counts)
                cap = min(int(max(most) * ghost_headroom) + 127 & ~127, min(int(c[2]) for c in counts))
                fixed = plan = FixedGhostPlan(comm, no, cap, g, dev)
                fixed.refresh(window[:, -1, 0])
                win = None                                             # (re)allocate the local system for this capacity
        else:
            plan = fixed
        own_win = window if win is None else win[nl:nl + no, :s].contiguous()
        gl, gr = exchange_rows(comm, plan, own_win)                    # full windows of the ghosts
        nl, nr, nloc = plan.n_left, plan.n_right, plan.n_local
        if win is None or fixed is None:
            win = torch.empty(nloc, s + 1, dim, dtype=torch.float32, device=dev)
            win[nl:nl + no, :s] = own_win
            shift = torch.empty(nloc, s, dim, dtype=torch.float32, device=dev)
            types = torch.zeros(nloc, dtype=torch.int64, device=dev)
            ptr = torch.tensor([0, nloc], dtype=torch.int32, device=dev)
            st_buf = torch.empty(nloc, dtype=torch.float32, device=dev)
            out_pos = torch.empty(no, dim, dtype=torch.float32, device=dev)
            out_str = torch.empty(no, dtype=torch.float32, device=dev)
            graph = None
        win[:nl, :s], win[nl + no:, :s] = gl, gr
        if particle_types_owned is not None:                          # ghosts carry their owners' types
            tl, tr = exchange_rows(comm, plan, particle_types_owned.to(dev).long().contiguous())
            types[:nl], types[nl:nl + no], types[nl + no:] = tl, particle_types_owned.to(dev).long(), tr

        if k == 0:       # kernel plans and latent buffers with headroom: the local system may change size at a refresh
            simulator._reserve(nloc + max(nloc // 16, 1024), dev)

        def step():
            return simulator._step(win[:, :s], None, types, next_pos=win[:, s], strain=st_buf, example_ptr=ptr)

        def after_step():
            """owned results out, ghosts take their owners' new frame, window slides by one frame"""
            out_pos.copy_(win[nl:nl + no, s])
            out_str.copy_(st_buf[nl:nl + no])
            fl, fr = exchange_rows(comm, plan, out_pos)
            win[:nl, s] = fl
            win[nl + no:, s] = fr
            shift.copy_(win[:, 1:])                                    # frames 1..S -> 0..S-1
            win[:, :s] = shift

        # With fixed message sizes over NCCL the whole per-step body -- step, owned results, ghost-frame exchange,
        # window shift -- is ONE graph: no Python-issued launch between two steps.  (ThreadComm synchronises on the
        # host and a host-planned refresh changes sizes: those keep the exchange outside the graph.)
        whole = use_graph and fixed is not None and isinstance(comm, DistComm) and not os.environ.get("SGNN_SHARD_EAGER_EXCHANGE")
        for j in range(m):
            if use_graph and graph is None and (m - j > 1 or fixed is not None) and (j == 1 or (j == 0 and k > 0)):
                # the very first step of the rollout runs eagerly (lazy initialisation, the step-0 edge list);
                # every other step replays the captured step
                simulator._prepare(nloc, dev)                          # never build kernel plans inside a capture
                graph = torch.cuda.CUDAGraph()
                cap_stream = torch.cuda.Stream(device=dev)             # (torch.cuda.graph() would also empty the
                cap_stream.wait_stream(torch.cuda.current_stream())    #  allocator cache at every capture)
                with torch.cuda.stream(cap_stream):
                    graph.capture_begin(capture_error_mode="thread_local")
                    held = step()      # noqa: F841 -- the engine / edge set whose pointers the graph replays stay alive with it
                    if whole:
                        after_step()
                    graph.capture_end()
                torch.cuda.current_stream().wait_stream(cap_stream)
            if graph is not None:
                graph.replay()
                if not whole:
                    after_step()
            else:
                _, csr = step()
                if return_edges and edges0 is None:
                    ei = csr.edge_index()
                    keep = (ei[1] >= nl) & (ei[1] < nl + no)
                    edges0 = (ei[:, keep], nl)
                after_step()
            pred[k + j] = out_pos
            strain[k + j] = out_str
        if fixed is None:
            window = win[nl:nl + no, :s].contiguous()
            graph = None                                               # the local system changes size: re-capture
        k += m
    if return_edges:
        return pred, strain, edges0
    return pred, strain
