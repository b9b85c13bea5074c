#!/usr/bin/env python
"""bench.py -- semi-dense key frames/s of the B200-native ProbabilityMapping path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
           --master-port P bench.py --gpus N --steps K --warmup W

Workload (BASELINE.json configs[1]): 200 synthetic 1280x720 key frames (chris_logic_HD720
intrinsics), 10 covisible neighbours, per GPU.  With N > 1 the sequence is N*200 key frames,
sharded by contiguous index range, one rank per GPU, with one NCCL point-to-point exchange of the
boundary (inverse depth, sigma) maps between the intra- and inter-key-frame stages (weak scaling).

One step = one pass of the whole path over the rank's key frames:
    K0 gradient front end -> K1 epipolar search + fusion -> K2/K3 intra checks -> [halo exchange]
    -> K4 inter-key-frame check -> K5 point set.
`value`  : inputs resident in HBM (device u8 images), results stay on the device.
`e2e`    : the same through the C ABI with HOST buffers: pinned u8 images are copied host->device
           and the (inverse depth, sigma) maps device->host inside the timed region, every step.
Timing: CUDA events over all of the context's streams (sd_span_begin/end), barrier +
torch.cuda.synchronize() on both sides, max over ranks.  L2: the working set per step (~5 GB)
is far larger than the 126 MB L2, so no explicit flush is needed ("inputs_larger_than_l2").

`roofline` is quoted on stage K1 (k1_search + k1_fuse): `avg_launch_us` comes from the events of the timed region,
where the context's streams overlap on purpose (a kernel's events then also span the time it shares the SMs with
kernels of other key frames); `standalone_launch_us` / `kernels_standalone_us` come from one extra untimed pass with
every stream drained before each kernel -- those are the durations the ncu launch list (profiles/) agrees with.

--impl reference times the reference's CPU algorithm for the path (the oracle port, all host
threads) on a bounded sample of the same workload; rank 0 only.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
from __graft_entry__ import PKG_NAME, load_package  # noqa: E402

METRIC = "semi-dense keyframes/sec"
UNIT = "keyframes/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--camera", default="720p")
    ap.add_argument("--nkf", type=int, default=200, help="key frames per GPU")
    ap.add_argument("--nbr", type=int, default=10, help="covisible neighbours")
    ap.add_argument("--cpu-sample", type=int, default=12, help="key frames in the CPU-baseline sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--seed", type=int, default=1234)
    return ap.parse_args()


def algorithmic_bytes_per_px(n):
    """K0 9, K1 17+9N, K2 16, K3 20, K4 8+8N, K5 16."""
    return {"grad": 9, "search": 17 + 9 * n, "intra": 16, "grow": 20, "inter": 8 + 8 * n, "points": 16}


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md recipe)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index):
        self.idx, self.p, self.lines = gpu_index, None, []

    def start(self):
        try:
            self.p = subprocess.Popen(["nvidia-smi", "-i", str(self.idx), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits", "-lms", "100"],
                                      stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=lambda: [self.lines.append(l) for l in self.p.stdout], daemon=True)
            self.t.start()
        except Exception:
            self.p = None

    def stop(self):
        if not self.p:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.p.terminate()
        try:
            self.p.wait(timeout=5)
        except Exception:
            self.p.kill()
        sm, mx, pw, reasons = [], [], [], set()
        for l in self.lines:
            f = [x.strip() for x in l.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the oracle port on the host cores
# ------------------------------------------------------------------------------------------------
def cpu_pass(po, synth, pipeline, args, sample_kf, images=None, plan=None):
    """One pass of the whole path (front end .. point set) over `sample_kf` key frames with the
    CPU oracle on all host threads; returns (seconds, plan, images)."""
    if plan is None:
        scene = synth.Scene(args.camera, seed=args.seed)
        plan = pipeline.SequencePlan(scene, sample_kf, min(args.nbr, sample_kf - 1))
        images = {k: scene.render(k) for k in plan.ids}
    threads = po.lib().ora_max_threads()
    t0 = time.perf_counter()
    ora = po.Oracle(threads=threads)
    for k in plan.ids:
        ora.add_keyframe(k, images[k], plan.scene.K, plan.poses[k])
    for k in plan.ids:
        ora.semidense(k, plan.nbrs[k], plan.rots[k], *plan.prior[k])
    for k in plan.ids:
        ora.interkf(k, plan.nbrs[k])
        ora.pointset(k)
    dt = time.perf_counter() - t0
    ora.close()
    return dt, plan, images, threads


def workload_name(args, world):
    return (f"{args.nkf} synthetic {args.camera} key frames per GPU (chris_logic_HD720 intrinsics), "
            f"{args.nbr} covisible neighbours" + (f", {world} ranks keyframe-sharded with NCCL map exchange" if world > 1 else ""))


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    load_package()
    synth = importlib.import_module(PKG_NAME + ".synth")
    pipeline = importlib.import_module(PKG_NAME + ".pipeline")
    from oracle import pm_oracle as po
    sample = max(args.cpu_sample, args.nbr + 1)
    _, plan, images, threads = cpu_pass(po, synth, pipeline, args, sample)   # also warms caches / builds
    for _ in range(max(args.warmup - 1, 0)):
        cpu_pass(po, synth, pipeline, args, sample, images, plan)
    t = 0.0
    for _ in range(args.steps):
        t += cpu_pass(po, synth, pipeline, args, sample, images, plan)[0]
    kfps = sample * args.steps / t
    line = {
        "impl": "reference", "metric": METRIC, "value": kfps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(args, 1), "camera": args.camera, "neighbours": args.nbr},
        "cpu_baseline": {"value": kfps, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{sample} of the {args.nkf} key frames per step, all stages K0..K5, OpenMP on the reference's five pragma loops"},
        "e2e": {"value": kfps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "pixels_per_s": kfps * plan.scene.w * plan.scene.h,
        "note": "reference CPU algorithm = oracle/pm_oracle.cc (PM.cc is not compilable: SURVEY F1-F3); per-call cv::Mat churn hoisted, so faster than the literal reference",
    }
    emit(line)


# ------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------
def bind_to_gpu_numa_node(torch, local):
    """One process per GPU on a multi-socket box: run the rank (and first-touch its pinned staging buffers) on the
    CPUs of the GPU's own NUMA node, so that the host<->device copies of the e2e leg do not cross the socket link.
    Best effort: returns the cpu list string it bound to, or None."""
    try:
        pr = torch.cuda.get_device_properties(local)
        bus = "%04x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
        path = "/sys/bus/pci/devices/%s/local_cpulist" % bus.lower()
        with open(path) as f:
            cpulist = f.read().strip()
        cpus = set()
        for part in cpulist.split(","):
            if "-" in part:
                a, b = part.split("-")
                cpus.update(range(int(a), int(b) + 1))
            elif part:
                cpus.add(int(part))
        cpus &= os.sched_getaffinity(0)
        if len(cpus) >= 2:
            os.sched_setaffinity(0, cpus)
            return cpulist
    except Exception:
        pass
    return None



def run_b200(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa = bind_to_gpu_numa_node(torch, local) if world > 1 else None   # before any pinned allocation (first touch)

    pkg = load_package()
    synth = importlib.import_module(PKG_NAME + ".synth")
    pipeline = importlib.import_module(PKG_NAME + ".pipeline")
    sharding = importlib.import_module(PKG_NAME + ".sharding")

    n = args.nbr
    total_kf = args.nkf * world
    scene = synth.Scene(args.camera, seed=args.seed)
    W, H = scene.w, scene.h
    gplan = pipeline.SequencePlan(scene, total_kf, n) if world == 1 else None
    nbr_of = (lambda k: gplan.nbrs[k]) if gplan else (lambda k: synth.Scene.neighbours(k, total_kf, n))
    shard = sharding.make_plan(total_kf, world, rank, nbr_of)
    owned, resident = shard.owned, shard.resident
    if gplan is None:  # host-side inputs for the owned key frames only
        class P:  # minimal stand-in for SequencePlan restricted to this rank
            pass
        gplan = P()
        gplan.scene = scene
        gplan.poses = {k: scene.pose(k) for k in resident}
        gplan.nbrs = {k: nbr_of(k).astype(np.int32) for k in owned}
        gplan.rots = {k: np.array([synth.Scene.rot_in_plane(k, j) for j in gplan.nbrs[k]], np.float32) for k in owned}
        gplan.prior = {k: pkg.stereo_search_constraints(scene.sparse_depths(k, 1000)) for k in owned}

    # inputs: pinned host images (e2e leg) and a device-resident copy (value leg)
    host_img = torch.empty((len(resident), H, W), dtype=torch.uint8).pin_memory()
    hnp = host_img.numpy()
    for i, k in enumerate(resident):
        scene.render(k, out=hnp[i])
    dev_img = host_img.to(dev)
    slot = {k: i for i, k in enumerate(resident)}
    host_out = torch.empty((len(owned), 2, H, W), dtype=torch.float32).pin_memory()
    oslot = {k: i for i, k in enumerate(owned)}
    img_bytes = W * H

    ctx = pkg.Context(local)
    ext = torch.cuda.ExternalStream(ctx.compute_stream(), device=dev)
    send_buf = {p: torch.empty((len(ids), 2, H, W), dtype=torch.float32, device=dev) for p, ids in shard.send.items()}
    recv_buf = {p: torch.empty((len(ids), 2, H, W), dtype=torch.float32, device=dev) for p, ids in shard.recv.items()}

    def pack(ids, peer):
        t = send_buf[peer]
        for i, k in enumerate(ids):
            ctx.export_maps(k, t[i, 0].data_ptr(), t[i, 1].data_ptr())
        return t

    def exchange():
        if world == 1:
            return 0
        if os.environ.get("BENCH_SKIP_NCCL"):   # diagnostic: same packing / installing, no wire (results are then wrong at the seams)
            for p, ids in shard.send.items():
                pack(ids, p)
            for p, ids in shard.recv.items():
                for i, k in enumerate(ids):
                    ctx.import_maps(k, recv_buf[p][i, 0].data_ptr(), recv_buf[p][i, 1].data_ptr())
            return 0
        with torch.cuda.stream(ext):  # NCCL orders against the context's compute stream
            ops = []
            for p in sorted(set(shard.send) | set(shard.recv)):
                if p in shard.send:
                    ops.append(dist.P2POp(dist.isend, pack(shard.send[p], p), p))
                if p in shard.recv:
                    ops.append(dist.P2POp(dist.irecv, recv_buf[p], p))
            for r in dist.batch_isend_irecv(ops):
                r.wait()
            for p, ids in shard.recv.items():
                for i, k in enumerate(ids):
                    ctx.import_maps(k, recv_buf[p][i, 0].data_ptr(), recv_buf[p][i, 1].data_ptr())
        return sum(t.numel() * 4 for t in list(send_buf.values()) + list(recv_buf.values()))

    last_needed = {k: int(max(gplan.nbrs[k].max(), k)) for k in owned}
    schedule = sharding.pipelined_schedule(shard, lambda k: gplan.nbrs[k]) if world > 1 else None
    plain_schedule = [("semidense", k) for k in owned] + [("exchange",)] + [("finish", k) for k in owned]

    def step(e2e):
        for k in resident:
            if e2e:
                ctx.upload_keyframe_ptr(k, host_img[slot[k]].data_ptr(), W, H, W, scene.K, gplan.poses[k])
            else:
                ctx.upload_keyframe_ptr(k, dev_img[slot[k]].data_ptr(), W, H, W, scene.K, gplan.poses[k], device=True)

        def finish(j):
            ctx.interkf_check(j, gplan.nbrs[j])
            ctx.update_pointset(j)
            if e2e:
                ctx.download_async_ptr(j, host_out[oslot[j], 0].data_ptr(), host_out[oslot[j], 1].data_ptr())

        if world == 1:
            # software pipeline: the inter-KF check of key frame j is issued as soon as every neighbour
            # of j has its intra-KF maps queued, so downloads overlap the remaining search kernels
            pending = list(owned)
            for k in owned:
                mn, mx = gplan.prior[k]
                ctx.semidense(k, gplan.nbrs[k], gplan.rots[k], mn, mx)
                while pending and last_needed[pending[0]] <= k:
                    finish(pending.pop(0))
            for j in pending:
                finish(j)
        else:
            # boundary key frames first, one exchange, then the interior with the same software pipeline
            # (sharding.pipelined_schedule): only the few key frames with halo neighbours wait for the peers
            for op in (plain_schedule if os.environ.get("BENCH_PLAIN_SCHEDULE") else schedule):
                if op[0] == "semidense":
                    k = op[1]
                    mn, mx = gplan.prior[k]
                    ctx.semidense(k, gplan.nbrs[k], gplan.rots[k], mn, mx)
                elif op[0] == "exchange":
                    exchange()
                else:
                    finish(op[1])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    host_issue_ms = []

    def timed(e2e, steps, warmup, sampler=None):
        for _ in range(warmup):
            step(e2e)
            ctx.sync()
        barrier()
        if sampler:
            sampler.start()
        ctx.timing_reset()
        l0 = ctx.launch_count()
        ctx.span_begin()
        t_host = time.perf_counter()
        for _ in range(steps):
            step(e2e)
        host_issue_ms.append(1e3 * (time.perf_counter() - t_host))   # host time to ISSUE the steps (no device wait)
        ms = ctx.span_end()
        barrier()
        clocks = sampler.stop() if sampler else None
        launches = ctx.launch_count() - l0
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if os.environ.get("BENCH_VERBOSE"):
            sys.stderr.write(f"rank {rank}: numa cpus {numa}; span {ms:.2f} ms for {steps} steps, host issue {host_issue_ms[-1]:.2f} ms\n")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), launches, clocks

    ctx.timing_enable(True)
    sampler = ClockSampler(local) if rank == 0 else None
    ms_v, launches, clocks = timed(False, args.steps, args.warmup, sampler)
    kt = ctx.timing()
    ctx.timing_enable(False)
    ms_e, _, _ = timed(True, args.steps, max(args.warmup, 1))

    # standalone kernel durations: one more untimed pass with every stream drained before each kernel, so a launch's
    # events bracket that kernel alone.  In the timed region the five streams overlap on purpose; there the events
    # of a kernel also span the time it shares the SMs with (or waits behind) kernels of other key frames.
    ctx.timing_enable(True, serial=True)
    ctx.timing_reset()
    step(False)
    ctx.sync()
    kt_alone = ctx.timing()
    ctx.timing_enable(False)

    kf_total_per_step = args.nkf * world
    value = kf_total_per_step * args.steps / (ms_v * 1e-3)
    e2e = kf_total_per_step * args.steps / (ms_e * 1e-3)

    # sanity: the maps that came back are real results
    d = host_out[len(owned) // 2, 0].numpy()
    valid_fraction = float((d > 1e-6).mean())

    if rank == 0:
        peak, peak_src = measured_peak()
        px = W * H
        bpp = algorithmic_bytes_per_px(n)
        k1_ms, k1_n = kt["search"]          # stage K1 = k1_search + k1_fuse (one launch each per key frame)
        k1_ms += kt["fuse"][0]
        k1_avg_s = (k1_ms / max(k1_n, 1)) * 1e-3
        achieved = bpp["search"] * px / k1_avg_s / 1e9 if k1_avg_s > 0 else 0.0
        alone_us = {k: (v[0] / max(v[1], 1)) * 1e3 for k, v in kt_alone.items()}
        k1_alone_s = (alone_us["search"] + alone_us["fuse"]) * 1e-6
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "k1_traffic.json")) as f:
                traffic = json.load(f).get("dram_bytes_per_launch")
        except Exception:
            pass
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_v / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": workload_name(args, world), "camera": args.camera, "width": W, "height": H,
                       "neighbours": n, "keyframes_per_gpu": args.nkf, "l2": "inputs_larger_than_l2",
                       "parallelism": f"keyframe-shard x{world}" if world > 1 else "single GPU",
                       "e2e_result": "inverse-depth + sigma maps to pinned host memory; point sets stay device-resident"},
            "pixels_per_s": value * px,
            "e2e": {"value": e2e, "unit": UNIT, "ms_per_step": ms_e / args.steps,
                    "h2d_bytes_per_step": len(resident) * img_bytes, "d2h_bytes_per_step": len(owned) * 2 * px * 4},
            "gpu_launches": int(launches),
            "host_issue_ms_per_step": {"value": host_issue_ms[0] / args.steps, "e2e": host_issue_ms[1] / args.steps},
            "clocks": clocks,
            "roofline": {"kernel": "k1_search + k1_fuse (epipolar search + hypothesis fusion)", "bound": "hbm", "achieved": achieved,
                         "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "algorithmic_bytes_per_launch": bpp["search"] * px,
                         "avg_launch_us": k1_avg_s * 1e6,
                         "standalone_launch_us": k1_alone_s * 1e6,
                         "standalone_achieved": bpp["search"] * px / k1_alone_s / 1e9 if k1_alone_s > 0 else 0.0,
                         "whole_path_frac": sum(bpp.values()) * px * (value / world) / 1e9 / peak},
            "kernels_avg_us": {k: (v[0] / max(v[1], 1)) * 1e3 for k, v in kt.items()},
            "kernels_standalone_us": alone_us,
            "kernels_standalone_share": {k: v / max(sum(alone_us.values()), 1e-9) for k, v in alone_us.items()},
            "valid_fraction": valid_fraction,
        }
        if world == 1 and not args.no_cpu_baseline:
            from oracle import pm_oracle as po
            sample = max(args.cpu_sample, n + 1)
            dt, _, _, threads = cpu_pass(po, synth, pipeline, args, sample)
            line["cpu_baseline"] = {"value": sample / dt, "unit": UNIT, "cores": threads, "kind": "port",
                                    "sample": f"first {sample} key frames of the same workload, all stages K0..K5, oracle port with OpenMP on the reference's five pragma loops (one pass, {dt:.1f} s)"}
        emit(line)
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


_JSON_FD = None


def emit(line):
    """The ONE JSON line of the contract, on the real stdout."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def main():
    global _JSON_FD
    args = parse()
    # libraries print to stdout behind our back (NCCL: "NCCL version ..." at communicator creation): keep fd 1 for
    # the JSON line only, everything else goes to stderr
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
