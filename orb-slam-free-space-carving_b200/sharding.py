"""Key-frame sharding across the GPUs of one box.

K0..K3 of a key frame read only IMAGE planes of its neighbours; K4 (InterKeyFrameDepthChecking,
PM.cc:628-799) reads the neighbours' post-intra (inverse depth, sigma) maps.  So the sequence is
cut into contiguous index ranges, one per rank (one process per GPU), and there is exactly one
exchange step: after K3 every rank sends the maps of the boundary key frames its neighbours need
and receives the ones it needs.  On the GPUs that step is sd_halo_exchange behind the C ABI (grouped
ncclSend / ncclRecv straight out of / into the key-frame planes, include/sdpm.h); the CPU tests run
the same plan with gloo point-to-point ops and host buffers (exchange_maps below).  Halo key frames' images are uploaded (and K0 run) redundantly on the
ranks that need them: recomputing 9 B/px is cheaper than shipping 8 B/px of planes.

Everything here is a pure function of (global key-frame count, world size, neighbour lists), so
every rank derives the same plan without communication.
"""
from dataclasses import dataclass, field
from typing import Callable, Dict, List

import numpy as np


def owner_ranges(nkf: int, world: int):
    """Contiguous, balanced ranges: rank r owns [start[r], start[r+1])."""
    base, rem = divmod(nkf, world)
    starts = [0]
    for r in range(world):
        starts.append(starts[-1] + base + (1 if r < rem else 0))
    return starts


def owner_of(k: int, starts) -> int:
    return int(np.searchsorted(np.asarray(starts[1:]), k, side="right"))


@dataclass
class ShardPlan:
    rank: int
    world: int
    owned: List[int]
    image_halo: List[int]                 # not owned, images needed for K1 of owned key frames
    recv: Dict[int, List[int]] = field(default_factory=dict)   # peer -> key frames whose maps we receive
    send: Dict[int, List[int]] = field(default_factory=dict)   # peer -> key frames whose maps we send

    @property
    def resident(self):
        return sorted(self.owned + self.image_halo)


def make_plan(nkf: int, world: int, rank: int, neighbours: Callable[[int], np.ndarray]) -> ShardPlan:
    """neighbours(k) -> ordered covisible key-frame ids of k (GetVectorCovisibleKeyFrames stand-in)."""
    starts = owner_ranges(nkf, world)

    def needs(r):
        own = range(starts[r], starts[r + 1])
        need = set()
        for k in own:
            need.update(int(j) for j in neighbours(k))
        return sorted(need - set(own))

    mine = needs(rank)
    plan = ShardPlan(rank, world, list(range(starts[rank], starts[rank + 1])), mine)
    for k in mine:
        plan.recv.setdefault(owner_of(k, starts), []).append(k)
    for peer in range(world):
        if peer == rank:
            continue
        theirs = [k for k in needs(peer) if starts[rank] <= k < starts[rank + 1]]
        if theirs:
            plan.send[peer] = theirs
    return plan


def exchange_maps(plan: ShardPlan, dist, pack: Callable[[List[int]], "object"], alloc: Callable[[int], "object"],
                  unpack: Callable[[List[int], "object"], None]):
    """One halo exchange.  pack(ids) -> tensor [len(ids), 2, H, W] of (inverse depth, sigma);
    alloc(n) -> empty receive tensor for n key frames; unpack(ids, tensor) installs received maps.
    Returns bytes sent + received by this rank."""
    ops, recv_bufs, keep = [], [], []
    nbytes = 0
    for peer in sorted(set(plan.send) | set(plan.recv)):
        if peer in plan.send:
            t = pack(plan.send[peer])
            keep.append(t)
            ops.append(dist.P2POp(dist.isend, t, peer))
            nbytes += t.numel() * t.element_size()
        if peer in plan.recv:
            t = alloc(len(plan.recv[peer]))
            recv_bufs.append((plan.recv[peer], t))
            ops.append(dist.P2POp(dist.irecv, t, peer))
            nbytes += t.numel() * t.element_size()
    if ops:
        for req in dist.batch_isend_irecv(ops):
            req.wait()
    for ids, t in recv_bufs:
        unpack(ids, t)
    return nbytes


def pipelined_schedule(plan: ShardPlan, neighbours: Callable[[int], np.ndarray]):
    """Issue order of one rank's step that keeps the stream pipeline of the single-GPU path
    (search | fuse+intra | inter+points) running across the halo exchange:

        1. ("semidense", k) for the BOUNDARY key frames (the ones some peer needs), so their maps exist early;
        2. ("exchange",)    the one point-to-point exchange step (send boundary maps, receive + install halo maps);
        3. ("semidense", k) for the interior key frames, each followed by ("finish", j) -- inter-key-frame
           check + point set -- for every owned j whose neighbours all have maps by then.

    A "finish" is emitted only after the semidense of every OWNED neighbour and, for key frames with halo
    neighbours, after the exchange.  Pure function of the plan: every rank derives it locally.
    """
    owned = list(plan.owned)
    own = set(owned)
    boundary = sorted({k for ids in plan.send.values() for k in ids})
    bset = set(boundary)
    order = boundary + [k for k in owned if k not in bset]
    nb = {j: [int(v) for v in neighbours(j)] for j in owned}
    waiting = {j: sum(1 for v in nb[j] if v in own) for j in owned}   # owned neighbours without maps yet
    users = {}
    for j in owned:
        for v in nb[j]:
            if v in own:
                users.setdefault(v, []).append(j)
    has_halo = {j: any(v not in own for v in nb[j]) for j in owned}
    done_self = set()
    ops, ready_wait_exchange = [], []
    exchanged = not (plan.send or plan.recv)

    def maybe_finish(j):
        if waiting[j] == 0 and j in done_self:
            if has_halo[j] and not exchanged:
                ready_wait_exchange.append(j)
            else:
                ops.append(("finish", j))

    for i, k in enumerate(order):
        if i == len(boundary) and not exchanged:
            ops.append(("exchange",))
            exchanged = True
            for j in ready_wait_exchange:
                ops.append(("finish", j))
            ready_wait_exchange = []
        ops.append(("semidense", k))
        done_self.add(k)
        for j in users.get(k, []):
            waiting[j] -= 1
            if j != k:
                maybe_finish(j)
        maybe_finish(k)
    if not exchanged:      # every owned key frame is a boundary key frame
        ops.append(("exchange",))
        for j in ready_wait_exchange:
            ops.append(("finish", j))
    return ops
