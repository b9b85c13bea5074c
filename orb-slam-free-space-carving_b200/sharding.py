"""Key-frame sharding across the GPUs of one box.

K0..K3 of a key frame read only IMAGE planes of its neighbours; K4 (InterKeyFrameDepthChecking,
PM.cc:628-799) reads the neighbours' post-intra (inverse depth, sigma) maps.  So the sequence is
cut into contiguous index ranges, one per rank (one process per GPU), and there is exactly one
exchange step: after K3 every rank sends the maps of the boundary key frames its neighbours need
and receives the ones it needs, as point-to-point torch.distributed ops (NCCL over NVLink on the
GPUs, gloo in the CPU tests).  Halo key frames' images are uploaded (and K0 run) redundantly on the
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
