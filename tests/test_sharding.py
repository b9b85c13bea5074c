"""Multi-GPU host logic on the CPU: the shard plan, and the one halo exchange of the path
 run for real with torch.distributed `gloo`, world_size 2 and 3."""
import importlib
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp


def _mods(built):
    built.load_package()
    return (importlib.import_module(built.PKG_NAME + ".sharding"), importlib.import_module(built.PKG_NAME + ".synth"))


@pytest.mark.parametrize("nkf,world,n", [(400, 2, 10), (1000, 8, 20), (37, 4, 6), (12, 2, 10), (50, 3, 40)])
def test_plan_is_consistent(built, nkf, world, n):
    sharding, synth = _mods(built)
    nb = lambda k: synth.Scene.neighbours(k, nkf, n)
    plans = [sharding.make_plan(nkf, world, r, nb) for r in range(world)]
    owned = sorted(k for p in plans for k in p.owned)
    assert owned == list(range(nkf))                                   # a partition
    assert max(len(p.owned) for p in plans) - min(len(p.owned) for p in plans) <= 1
    for p in plans:
        assert all(b == a + 1 for a, b in zip(p.owned, p.owned[1:]))  # contiguous
        need = set()
        for k in p.owned:
            need.update(int(j) for j in nb(k))
        assert set(p.image_halo) == need - set(p.owned)               # exactly what K1/K4 read
        assert sorted(k for ids in p.recv.values() for k in ids) == p.image_halo
        for peer, ids in p.recv.items():
            assert plans[peer].send[p.rank] == ids                    # mirror image on the sender
            assert set(ids) <= set(plans[peer].owned)
        for peer, ids in p.send.items():
            assert plans[peer].recv[p.rank] == ids


@pytest.mark.parametrize("nkf,world,n", [(400, 2, 10), (1000, 8, 20), (37, 4, 6), (12, 2, 10), (50, 3, 40), (25, 1, 10)])
def test_pipelined_schedule_respects_dependencies(built, nkf, world, n):
    """Issue order of a rank's step (bench.py, N > 1): every owned key frame gets exactly one semidense and one
    finish (inter-KF check + point set); the exchange comes after the semidense of everything this rank sends;
    a finish comes after the semidense of all owned neighbours and, with halo neighbours, after the exchange;
    interior finishes are interleaved with the remaining semidense calls (the pipeline keeps running)."""
    sharding, synth = _mods(built)
    nb = lambda k: synth.Scene.neighbours(k, nkf, n)
    for r in range(world):
        p = sharding.make_plan(nkf, world, r, nb)
        ops = sharding.pipelined_schedule(p, nb)
        own, sd, fin, exchanged = set(p.owned), [], [], False
        early = 0
        for op in ops:
            if op[0] == "semidense":
                sd.append(op[1])
            elif op[0] == "exchange":
                assert not exchanged
                exchanged = True
                assert all(k in sd for ids in p.send.values() for k in ids)
            else:
                j = op[1]
                fin.append(j)
                assert j in sd
                for v in (int(v) for v in nb(j)):
                    assert (v in sd) if v in own else exchanged, (r, j, v)
                early += len(sd) < len(p.owned)
        assert sorted(sd) == p.owned and sorted(fin) == p.owned
        assert exchanged == (world > 1)
        if len(p.owned) > 3 * n:
            assert early > len(p.owned) // 2          # most checks are queued while searches are still being issued


def test_interior_halo_size(built):
    sharding, synth = _mods(built)
    nb = lambda k: synth.Scene.neighbours(k, 1000, 20)
    p = sharding.make_plan(1000, 8, 3, nb)
    assert len(p.owned) == 125 and len(p.image_halo) == 20            # N/2 on each side
    assert p.recv[2] == list(range(365, 375)) and p.recv[4] == list(range(500, 510))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, nkf, n, root):
    import sys
    sys.path.insert(0, root)
    import __graft_entry__ as ge
    ge.load_package()
    sharding = importlib.import_module(ge.PKG_NAME + ".sharding")
    synth = importlib.import_module(ge.PKG_NAME + ".synth")
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    plan = sharding.make_plan(nkf, world, rank, lambda k: synth.Scene.neighbours(k, nkf, n))
    H, W = 6, 8
    maps = {k: torch.stack([torch.full((H, W), float(k)), torch.full((H, W), k + 0.5)]) for k in plan.owned}

    def pack(ids):
        return torch.stack([maps[k] for k in ids])

    def unpack(ids, t):
        for i, k in enumerate(ids):
            maps[k] = t[i].clone()

    nbytes = sharding.exchange_maps(plan, dist, pack, lambda m: torch.empty((m, 2, H, W)), unpack)
    assert nbytes == (sum(map(len, plan.send.values())) + sum(map(len, plan.recv.values()))) * 2 * H * W * 4
    for k in plan.image_halo:                                          # every halo map arrived, and is the right one
        assert float(maps[k][0, 0, 0]) == float(k) and float(maps[k][1, 0, 0]) == k + 0.5
    for k in plan.owned:                                               # now every owned key frame can run K4
        for j in synth.Scene.neighbours(k, nkf, n):
            assert int(j) in maps
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,nkf,n", [(2, 24, 6), (3, 31, 10)])
def test_halo_exchange_gloo(built, world, nkf, n):
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    mp.spawn(_worker, args=(world, _free_port(), nkf, n, root), nprocs=world, join=True)
