"""One rank of the key-frame-sharded path, used by tests/test_gpu_sharded.py and by
tools/gpu_shard_nccl.sh.  Launched once per rank with RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT set.

  SHARD_BACKEND=nccl : one GPU per rank; the (inverse depth, sigma) halo maps travel GPU->GPU through the C ABI's own
                       exchange (sd_comm_init + sd_halo_exchange: grouped ncclSend / ncclRecv straight out of / into
                       the key-frame planes); torch.distributed (gloo) only distributes the ncclUniqueId
  SHARD_BACKEND=gloo : every rank on cuda:0 (a single-GPU box), the same maps are staged through host
                       memory and travel over gloo -- same sharding plan, sd_export_maps / sd_import_maps,
                       same kernels

Every rank checks the key frames it owns against the CPU oracle run on the WHOLE sequence: sharding
must not change a single value.  Exit code 0 = parity."""
import importlib
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import PKG_NAME, load_package  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    backend = os.environ.get("SHARD_BACKEND", "gloo")
    cam = os.environ.get("SHARD_CAM", "small")
    nkf, n = int(os.environ.get("SHARD_NKF", "14")), int(os.environ.get("SHARD_NBR", "6"))
    pkg = load_package()
    synth = importlib.import_module(PKG_NAME + ".synth")
    pipeline = importlib.import_module(PKG_NAME + ".pipeline")
    sharding = importlib.import_module(PKG_NAME + ".sharding")
    from oracle import pm_oracle as po

    local = rank if backend == "nccl" else 0
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("gloo", rank=rank, world_size=world)   # plumbing only: unique id, staged maps, barrier

    small = dict(w=320, h=240, fx=262.5, fy=262.5, cx=159.5, cy=119.5)
    scene = synth.Scene(small if cam == "small" else cam, seed=77)
    H, W = scene.h, scene.w
    plan = pipeline.SequencePlan(scene, nkf, n)          # the whole sequence (host-side inputs are cheap)
    shard = sharding.make_plan(nkf, world, rank, lambda k: plan.nbrs[k])
    images = {k: scene.render(k) for k in plan.ids}

    ctx = pkg.Context(local)
    for k in shard.resident:                              # owned + image halo: K0 is recomputed locally
        ctx.upload_keyframe(k, images[k], scene.K, plan.poses[k])
    ext = torch.cuda.ExternalStream(ctx.compute_stream(), device=dev)
    halo = None
    if backend == "nccl":
        ids = [pkg.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0)
        ctx.comm_init(world, rank, ids[0])
        halo = ctx.make_halo(shard.send, shard.recv)

    def pack(ids):
        t = torch.empty((len(ids), 2, H, W), dtype=torch.float32, device=dev)
        with torch.cuda.stream(ext):
            for i, k in enumerate(ids):
                ctx.export_maps(k, t[i, 0].data_ptr(), t[i, 1].data_ptr())
        ctx.sync()
        return t.cpu()

    def alloc(cnt):
        return torch.empty((cnt, 2, H, W), dtype=torch.float32, device="cpu")

    keep = []

    def unpack(ids, t):
        t = t.to(dev)
        keep.append(t)
        for i, k in enumerate(ids):
            ctx.import_maps(k, t[i, 0].data_ptr(), t[i, 1].data_ptr())

    # the issue order bench.py uses: boundary key frames, the one exchange, then the interior with the
    # inter-key-frame check of a key frame queued as soon as all its neighbours have maps
    nbytes = 0
    for op in sharding.pipelined_schedule(shard, lambda k: plan.nbrs[k]):
        if op[0] == "semidense":
            k = op[1]
            ctx.semidense(k, plan.nbrs[k], plan.rots[k], *plan.prior[k])
        elif op[0] == "exchange":
            if backend == "nccl":
                ctx.halo_exchange(halo)                   # on the context's compute stream, no staging
                nbytes = 2 * 4 * W * H * (sum(map(len, shard.send.values())) + sum(map(len, shard.recv.values())))
            else:
                with torch.cuda.stream(ext):
                    nbytes = sharding.exchange_maps(shard, dist, pack, alloc, unpack)
        else:
            ctx.interkf_check_points(op[1], plan.nbrs[op[1]])
    ctx.sync()

    ora = po.Oracle(threads=max(1, po.lib().ora_max_threads() // world))
    for k in plan.ids:
        ora.add_keyframe(k, images[k], scene.K, plan.poses[k])
    for k in plan.ids:
        ora.semidense(k, plan.nbrs[k], plan.rots[k], *plan.prior[k])
    bad = 0
    for k in shard.owned:
        ora.interkf(k, plan.nbrs[k])
        ora.pointset(k)
        for which in (po.DEPTH, po.SIGMA, po.DEPTH_CHECKED, po.POINTS):
            if not np.array_equal(ctx.plane(k, which), ora.plane(k, which)):
                bad += 1
                print(f"rank {rank}: key frame {k} plane {which} differs", flush=True)
    valid = sum(int((ctx.plane(k, po.DEPTH_CHECKED) > 1e-6).sum()) for k in shard.owned)
    print(f"rank {rank}/{world} [{backend}]: owned {shard.owned[0]}..{shard.owned[-1]}, image halo {shard.image_halo}, "
          f"exchanged {nbytes / 1e6:.1f} MB, valid px {valid}, mismatching planes {bad}", flush=True)
    ctx.close()
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(1 if bad or valid == 0 else 0)


if __name__ == "__main__":
    main()
