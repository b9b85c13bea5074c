"""Developer tool (GPU box): the box's device->host ceiling for the e2e leg -- every rank copies `count` buffers of
`mb` MB from its GPU to pinned host memory on one stream, no kernels; aggregate GB/s over ranks (max-over-ranks time).
Launch with torchrun at 1/2/4/8 ranks.  One JSON line per (world, size) on stdout of rank 0.
usage: torchrun --nproc-per-node N tools/d2h_ceiling.py [count]"""
import json
import os
import sys

import torch
import torch.distributed as dist


def main():
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    count = int(sys.argv[1]) if len(sys.argv) > 1 else 200
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    # 7.37 MB = both dense maps of a 720p key frame; 2.85 MB = its compact records; 5.7 MB = records with points
    for mb, what in ((7.3728, "dense depth+sigma planes"), (2.85, "compact records"), (5.7, "compact records + points")):
        nbytes = int(mb * 1e6)
        src = torch.empty(nbytes, dtype=torch.uint8, device=dev)
        dst = torch.empty((count, nbytes), dtype=torch.uint8).pin_memory()
        st = torch.cuda.Stream()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        best = None
        for rep in range(4):
            if world > 1:
                dist.barrier()
            torch.cuda.synchronize()
            with torch.cuda.stream(st):
                a.record(st)
                for i in range(count):
                    dst[i].copy_(src, non_blocking=True)
                b.record(st)
            torch.cuda.synchronize()
            ms = a.elapsed_time(b)
            t = torch.tensor([ms], dtype=torch.float64, device=dev)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            if rep > 0:
                best = min(best, float(t.item())) if best else float(t.item())
        if rank == 0:
            print(json.dumps({"ranks": world, "copy_mb": mb, "what": what, "copies_per_rank": count, "ms_max_over_ranks": best,
                              "gbs_per_rank": count * nbytes / best / 1e6, "gbs_aggregate": world * count * nbytes / best / 1e6,
                              "keyframes_per_s_ceiling": world * count / best * 1e3}), flush=True)
        del dst, src
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
