#!/usr/bin/env python
"""bench.py -- semi-dense key frames/s of the B200-native ProbabilityMapping path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
           --master-port P bench.py --gpus N --steps K --warmup W

Workload (BASELINE.json configs[1]): 200 synthetic 1280x720 key frames (chris_logic_HD720
intrinsics), 10 covisible neighbours, per GPU.  With N > 1 the sequence is N*200 key frames,
sharded by contiguous index range, one rank per GPU, with one NCCL point-to-point exchange of the
boundary (inverse depth, sigma) maps between the intra- and inter-key-frame stages (weak scaling);
the exchange is the C ABI's own (sd_halo_exchange: grouped ncclSend / ncclRecv on the key-frame planes).

One step = one pass of the whole path over the rank's key frames:
    K0 gradient front end -> K1 epipolar search + fusion -> K2/K3 intra checks -> [halo exchange]
    -> K4 inter-key-frame check + K5 point set (one launch).
`value`      : inputs resident in HBM (device u8 images), results stay on the device.
`e2e`        : the same through the C ABI with HOST buffers: pinned u8 images host->device and, device->host, the
               compact form of both result maps (one {pixel, inverse depth, sigma} record per pixel that carries
               anything; the consumer expands it: ProbabilityMapping::FetchResults) inside the timed region, every step.
`e2e_points` : e2e with the world point of every record as well (all three reference outputs, PM.cc:225-226, 361-363).
`e2e_dense`  : e2e with the two maps as dense planes (round 1's definition of e2e).
Timing: CUDA events over all of the context's streams (sd_span_begin/end), barrier +
torch.cuda.synchronize() on both sides, max over ranks.  L2: the working set per step (~5 GB)
is far larger than the 126 MB L2, so no explicit flush is needed ("inputs_larger_than_l2").
The timed legs run with per-kernel event timing OFF; `kernels_avg_us` / `roofline.avg_launch_us` come from one
extra pass of the same step with it on (the context's streams overlap on purpose: a kernel's events then also span the
time it shares the SMs with kernels of other key frames), `kernels_standalone_us` from one more pass with every stream
drained before each kernel -- those are the durations the ncu launch list (profiles/) agrees with.

With N > 1 every rank then checks its seam key frames (the owned key frames with neighbours on another rank) against the
CPU oracle, bitwise, and the line carries `seam_parity` (the only driver-visible proof of the NCCL path).

--impl reference times the reference's CPU algorithm for the path (the oracle port, all host
threads the process may use) on a bounded sample of the same workload; rank 0 only.
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
    ap.add_argument("--prior-sigmas", type=float, default=2.0, help="inverse-depth prior = mean +- s*sigma (PM.cc:381-382 uses 2)")
    ap.add_argument("--cpu-sample", type=int, default=20, help="key frames in the CPU-baseline sample (SURVEY 8d: a 20-key-frame slice)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-configs", action="store_true", help="skip the short runs of BASELINE configs 1, 3, 5")
    ap.add_argument("--no-seam-check", action="store_true")
    ap.add_argument("--no-class-leg", action="store_true", help="skip e2e_class (the C++ drop-in class driven by lib/test_pm_class --bench)")
    ap.add_argument("--class-window", type=int, default=32, help="key frames per sd_semidense_loop call of ProbabilityMapping::SemiDenseLoop")
    ap.add_argument("--class-threads", type=int, default=0, help="host threads that expand records into the KeyFrame members (0 = all cores, at most 16)")
    ap.add_argument("--seed", type=int, default=1234)
    return ap.parse_args()


def algorithmic_bytes_per_px(n):
    """K0 9, K1 17+9N, K2 16, K3 20, K4 8+8N, K5 16."""
    return {"grad": 9, "search": 17 + 9 * n, "intra": 16, "grow": 20, "inter": 8 + 8 * n, "points": 16}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            d = json.load(f)
        return float(d["hbm_gbs"]), "measured", float(d.get("sm_max_mhz", 1965.0))
    except Exception:
        return 6650.0, "fallback", 1965.0


def host_cores():
    """The CPUs this process may run on.  NOT omp_get_max_threads(): torchrun exports OMP_NUM_THREADS=1."""
    try:
        return max(1, len(os.sched_getaffinity(0)))
    except Exception:
        return max(1, os.cpu_count() or 1)


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


class NvmlSampler:
    """Per-rank clocks / power / throttle reasons of the rank's own GPU during the timed region, sampled in-process
    through NVML (what nvidia-smi reads): cheap enough for every rank of an 8-GPU run to do it.  Rank 0 also runs the
    recipe's nvidia-smi line (ClockSampler); this one is what `per_rank` reports."""

    def __init__(self, torch, local):
        self.h, self.nv, self.samples, self.stop_flag, self.t = None, None, [], False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            pr = torch.cuda.get_device_properties(local)
            bus = "%08x:%02x:%02x.0" % (pr.pci_domain_id, pr.pci_bus_id, pr.pci_device_id)
            self.h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
            self.nv = pynvml
        except Exception:
            self.h = None

    def _run(self):
        nv = self.nv
        while not self.stop_flag:
            try:
                self.samples.append((nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM),
                                     nv.nvmlDeviceGetPowerUsage(self.h) / 1000.0,
                                     nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)))
            except Exception:
                pass
            time.sleep(0.05)

    def start(self):
        if self.h is None:
            return
        self.samples, self.stop_flag = [], False
        self.t = threading.Thread(target=self._run, daemon=True)
        self.t.start()

    def stop(self):
        if self.h is None or self.t is None:
            return {"sm_mhz": None, "reasons": ["nvml unavailable"]}
        self.stop_flag = True
        self.t.join(timeout=2)
        nv, reasons = self.nv, set()
        names = {"hw_slowdown": "nvmlClocksEventReasonHwSlowdown", "hw_thermal_slowdown": "nvmlClocksEventReasonHwThermalSlowdown",
                 "sw_thermal_slowdown": "nvmlClocksEventReasonSwThermalSlowdown", "sw_power_cap": "nvmlClocksEventReasonSwPowerCap",
                 "hw_power_brake": "nvmlClocksEventReasonHwPowerBrakeSlowdown"}
        for _, _, mask in self.samples:
            for label, attr in names.items():
                bit = getattr(nv, attr, None) or getattr(nv, attr.replace("ClocksEventReason", "ClocksThrottleReason"), 0)
                if bit and (mask & bit):
                    reasons.add(label)
        sm = [x[0] for x in self.samples]
        pw = [x[1] for x in self.samples]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_mhz_min": float(min(sm)) if sm else None,
                "power_w_max": max(pw) if pw else None, "samples": len(sm), "reasons": sorted(reasons)}


def workload_name(args, world):
    return (f"{args.nkf} synthetic {args.camera} key frames per GPU (chris_logic_HD720 intrinsics), "
            f"{args.nbr} covisible neighbours" + (f", {world} ranks keyframe-sharded with NCCL map exchange" if world > 1 else ""))


def config_block(args, world, W, H, extra=None):
    """The keys both arms carry, so that the driver can tell they ran the same configuration."""
    c = {"workload": workload_name(args, world), "camera": args.camera, "width": W, "height": H,
         "neighbours": args.nbr, "keyframes_per_gpu": args.nkf, "prior_sigmas": args.prior_sigmas,
         "l2": "inputs_larger_than_l2",
         "parallelism": f"keyframe-shard x{world}" if world > 1 else "single GPU"}
    c.update(extra or {})
    return c


# ------------------------------------------------------------------------------------------------
# host-side inputs of a key-frame range (what ORB-SLAM's map would provide), without the product library:
# the reference arm and the seam check build them with the oracle's own StereoSearchConstraints
# ------------------------------------------------------------------------------------------------
class HostPlan:
    def __init__(self, synth, scene, total_kf, n, ids, prior_fn, prior_sigmas=2.0):
        self.scene, self.ids = scene, list(ids)
        self.poses, self.nbrs, self.rots, self.prior = {}, {}, {}, {}
        for k in self.ids:
            self.add(synth, total_kf, n, k, prior_fn, prior_sigmas)

    def add(self, synth, total_kf, n, k, prior_fn, prior_sigmas=2.0):
        if k in self.nbrs:
            return
        scene = self.scene
        self.poses[k] = scene.pose(k)
        nb = synth.Scene.neighbours(k, total_kf, n)
        self.nbrs[k] = nb.astype(np.int32)
        self.rots[k] = np.array([synth.Scene.rot_in_plane(k, int(j)) for j in nb], np.float32)
        d = scene.sparse_depths(k, 1000)
        if prior_sigmas == 2.0:
            self.prior[k] = prior_fn(d)                      # PM.cc:370-383 verbatim
        else:                                                # BASELINE config 5: widened prior
            mean, sd = float(np.mean(d)), float(np.std(d))
            self.prior[k] = (float(np.float32(1.0 / (mean - prior_sigmas * sd))), float(np.float32(1.0 / (mean + prior_sigmas * sd))))


# ------------------------------------------------------------------------------------------------
# reference arm / CPU baseline: the oracle port on the host cores
# ------------------------------------------------------------------------------------------------
def cpu_pass(po, plan, images, threads):
    """One pass of the whole path (front end .. point set) over the plan's key frames with the CPU oracle."""
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
    return dt


def cpu_sample_plan(args, synth, po, sample):
    scene = synth.Scene(args.camera, seed=args.seed)
    plan = HostPlan(synth, scene, sample, min(args.nbr, sample - 1), range(sample), po.stereo_search_constraints, args.prior_sigmas)
    images = {k: scene.render(k) for k in plan.ids}
    return scene, plan, images


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    load_package()                                   # module import only: the product library is never dlopen'ed in this arm
    synth = importlib.import_module(PKG_NAME + ".synth")
    from oracle import pm_oracle as po
    threads = host_cores()
    sample = max(args.cpu_sample, args.nbr + 1)
    scene, plan, images = cpu_sample_plan(args, synth, po, sample)
    for _ in range(max(args.warmup, 1)):
        cpu_pass(po, plan, images, threads)
    t = 0.0
    for _ in range(args.steps):
        t += cpu_pass(po, plan, images, threads)
    kfps = sample * args.steps / t
    # "as shipped": the reference builds without -fopenmp -> one thread; one pass over the smallest sample that
    # still has args.nbr neighbours per key frame
    s1 = args.nbr + 1
    _, plan1, images1 = cpu_sample_plan(args, synth, po, s1) if s1 != sample else (scene, plan, images)
    dt1 = cpu_pass(po, plan1, images1, 1)
    line = {
        "impl": "reference", "metric": METRIC, "value": kfps, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": config_block(args, max(args.gpus, 1), scene.w, scene.h),
        "sample_keyframes_per_step": sample,
        "cpu_baseline": {"value": kfps, "unit": UNIT, "cores": threads, "kind": "port",
                         "sample": f"{sample} of the {args.nkf} key frames per step, all stages K0..K5, OpenMP on the reference's five pragma loops"},
        "cpu_baseline_1thread": {"value": s1 / dt1, "unit": UNIT, "cores": 1, "kind": "port",
                                 "sample": f"{s1} key frames, one pass ({dt1:.1f} s): the reference as shipped builds without OpenMP (SURVEY F4)"},
        "e2e": {"value": kfps, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "pixels_per_s": kfps * scene.w * scene.h,
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


LEGS = ("value", "e2e", "e2e_points", "e2e_dense")


class Workload:
    """One rank's share of a key-frame sequence, resident inputs, pinned staging and the marshalled C calls of a step."""

    def __init__(self, torch, pkg, synth, sharding, ctx, dev, camera, nkf_per_rank, n, world, rank, seed, prior_sigmas=2.0,
                 legs=LEGS, record_fill=0.42):
        self.torch, self.pkg, self.ctx, self.world, self.rank, self.n = torch, pkg, ctx, world, rank, n
        self.total_kf = nkf_per_rank * world
        scene = self.scene = synth.Scene(camera, seed=seed)
        W, H = self.W, self.H = scene.w, scene.h
        nbr_of = lambda k: synth.Scene.neighbours(k, self.total_kf, n)
        shard = self.shard = sharding.make_plan(self.total_kf, world, rank, nbr_of)
        self.owned, self.resident = shard.owned, shard.resident
        self.plan = HostPlan(synth, scene, self.total_kf, n, self.owned, pkg.stereo_search_constraints, prior_sigmas)
        self.poses = {k: scene.pose(k) for k in self.resident}
        # inputs: pinned host images (e2e legs) and a device-resident copy (value leg)
        self.host_img = torch.empty((len(self.resident), H, W), dtype=torch.uint8).pin_memory()
        hnp = self.host_img.numpy()
        for i, k in enumerate(self.resident):
            scene.render(k, out=hnp[i])
        self.dev_img = self.host_img.to(dev)
        self.slot = {k: i for i, k in enumerate(self.resident)}
        # outputs: ONE pinned staging area, viewed per leg (dense planes / compact records)
        px = W * H
        self.cap = int(record_fill * px)              # records per key frame the staging area holds (about 26 % carry anything)
        per_kf = self.cap * pkg.COMPACT_DEPTH.itemsize                        # "e2e": compact records of the two maps
        if "e2e_points" in legs:
            per_kf = max(per_kf, self.cap * pkg.COMPACT_POINT.itemsize)
        if "e2e_dense" in legs:
            per_kf = max(per_kf, 2 * px * 4)
        self.out = torch.empty((len(self.owned), per_kf), dtype=torch.uint8).pin_memory()
        self.counts = torch.zeros((len(self.owned),), dtype=torch.int32).pin_memory()
        self.oslot = {k: i for i, k in enumerate(self.owned)}
        self.halo = ctx.make_halo(shard.send, shard.recv) if world > 1 else None
        # upload order = the order in which the step needs the planes: at world > 1 the boundary key frames are searched
        # first, so their images and their neighbours' go first (a rank whose boundary is at the END of its range would
        # otherwise run all its K0 launches alone before its first search, and its peer would wait for it at the exchange)
        self.upload_order = list(self.resident)
        if world > 1:
            first, seen = [], set()
            for k in sorted({k for ids in shard.send.values() for k in ids}):
                for j in [k] + [int(x) for x in self.plan.nbrs[k]]:
                    if j not in seen:
                        seen.add(j)
                        first.append(j)
            self.upload_order = first + [k for k in self.resident if k not in seen]
        self.loops = {leg: self._make_loops(leg) for leg in legs}

    def _download(self, leg, k):
        pkg, i = self.pkg, self.oslot[k]
        base = self.out[i].data_ptr()
        if leg == "value":
            return None
        if leg == "e2e":
            return (pkg.SD_DL_COMPACT, base, self.cap, self.counts[i:].data_ptr())
        if leg == "e2e_points":
            return (pkg.SD_DL_COMPACT_POINTS, base, self.cap, self.counts[i:].data_ptr())
        return (pkg.SD_DL_DENSE, base, base + self.W * self.H * 4)

    def _item(self, leg, k, finish_only=False, download=True):
        p = self.plan
        return dict(kf_id=k, nbrs=p.nbrs[k], rots=p.rots[k], prior=p.prior[k], finish_only=finish_only,
                    download=self._download(leg, k) if download else None)

    def _make_loops(self, leg):
        """world == 1: one sd_semidense_loop over the owned key frames.  world > 1: the boundary key frames first (the
        ones some peer needs), the exchange, then the boundary key frames' finishing stages + the interior."""
        ctx = self.ctx
        if self.world == 1:
            return [("loop", ctx.make_loop([self._item(leg, k) for k in self.owned]))]
        boundary = sorted({k for ids in self.shard.send.values() for k in ids})
        bset = set(boundary)
        interior = [k for k in self.owned if k not in bset]
        ops = [("loop", ctx.make_loop([self._item(leg, k, download=False) for k in boundary], interkf=False)),
               ("exchange", None),
               ("loop", ctx.make_loop([self._item(leg, k, finish_only=True) for k in boundary] + [self._item(leg, k) for k in interior]))]
        return ops

    def step(self, leg, mark=None):
        """`mark(label)`, if given, is called after the uploads and after every op of the step (timeline pass)."""
        ctx, W, H, K = self.ctx, self.W, self.H, self.scene.K
        src, on_dev = (self.dev_img, True) if leg == "value" else (self.host_img, False)
        for k in self.upload_order:
            ctx.upload_keyframe_ptr(k, src[self.slot[k]].data_ptr(), W, H, W, K, self.poses[k], device=on_dev)
        if mark:
            mark("uploads_issued")
        for op, arg in self.loops[leg]:
            if op == "loop":
                done = ctx.semidense_loop(arg)
                if arg["flags"] and not done.all():
                    raise RuntimeError("sd_semidense_loop skipped an inter-key-frame check")
            elif not os.environ.get("BENCH_SKIP_NCCL"):
                ctx.halo_exchange(self.halo)
            if mark:
                mark(op)

    def bytes_per_step(self, leg):
        px = self.W * self.H
        h2d = 0 if leg == "value" else len(self.resident) * px
        if leg == "value":
            d2h = 0
        elif leg == "e2e_dense":
            d2h = len(self.owned) * 2 * px * 4
        else:
            rec = self.pkg.COMPACT_POINT.itemsize if leg == "e2e_points" else self.pkg.COMPACT_DEPTH.itemsize
            d2h = int(self.counts.numpy().astype(np.int64).sum()) * rec + 4 * len(self.owned)
        return h2d, d2h

    def records(self, leg, k):
        i = self.oslot[k]
        n = int(self.counts[i])
        if n > self.cap:
            raise RuntimeError(f"key frame {k}: {n} records exceed the staging capacity {self.cap}")
        dt = self.pkg.COMPACT_POINT if leg == "e2e_points" else self.pkg.COMPACT_DEPTH
        return self.out[i].numpy()[:n * dt.itemsize].view(dt)


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
    sharding = importlib.import_module(PKG_NAME + ".sharding")
    if world > 1:
        # torchrun exports OMP_NUM_THREADS=1, which would render this rank's synthetic key frames (input generation,
        # outside every timed region) on one core: give the generator this rank's share of the host instead
        try:
            import ctypes
            ctypes.CDLL("libgomp.so.1").omp_set_num_threads(max(1, host_cores() // (1 if numa else world)))
        except Exception:
            pass

    ctx = pkg.Context(local)
    if world > 1:   # the path's own communicator: rank 0 draws the id, torch.distributed only carries the 128 bytes
        ids = [pkg.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0, device=dev)
        ctx.comm_init(world, rank, ids[0])
    wl = Workload(torch, pkg, synth, sharding, ctx, dev, args.camera, args.nkf, args.nbr, world, rank, args.seed, args.prior_sigmas)
    W, H, n = wl.W, wl.H, args.nbr

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    host_issue_ms = {}
    rank_ms = {}                       # this rank's own span per leg (the line's figures are the max over ranks)
    nvml = NvmlSampler(torch, local)
    rank_clocks = {}

    def timed(work, leg, steps, warmup, sampler=None):
        c = work.ctx
        for _ in range(warmup):
            work.step(leg)
            c.sync()
        barrier()
        if sampler:
            sampler.start()
        if leg in ("value", "e2e"):
            nvml.start()
        l0 = c.launch_count()
        c.span_begin()
        t_host = time.perf_counter()
        for _ in range(steps):
            work.step(leg)
        host_issue_ms[leg] = 1e3 * (time.perf_counter() - t_host) / steps   # host time inside the issue loop: includes waiting on a full launch queue / download ring (tools/stage_costs.py measures the unblocked cost)
        ms = c.span_end()
        if leg in ("value", "e2e"):
            rank_clocks[leg] = nvml.stop()
        barrier()
        clocks = sampler.stop() if sampler else None
        launches = c.launch_count() - l0
        rank_ms[leg] = ms / steps
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        if os.environ.get("BENCH_VERBOSE"):
            sys.stderr.write(f"rank {rank}: numa cpus {numa}; {leg}: span {ms:.2f} ms for {steps} steps, host issue {host_issue_ms[leg]:.2f} ms/step\n")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()), launches, clocks

    sampler = ClockSampler(local) if rank == 0 else None
    ms, launches, clocks, bytes_ps = {}, {}, None, {}
    for leg in LEGS:
        ms[leg], launches[leg], ck = timed(wl, leg, args.steps, args.warmup if leg == "value" else max(min(args.warmup, 2), 1),
                                           sampler if leg == "value" else None)
        clocks = clocks or ck
        bytes_ps[leg] = wl.bytes_per_step(leg)
        if leg in ("e2e", "e2e_points"):   # the records that came back are real results: expand one key frame, compare with the planes
            k = wl.owned[len(wl.owned) // 2]
            got = pkg.expand_compact(wl.records(leg, k), H, W, points=leg == "e2e_points")
            ref = ctx.download(k, points=leg == "e2e_points")
            if not all(np.array_equal(a, b) for a, b in zip(got, ref)):
                raise RuntimeError(f"{leg}: compact records of key frame {k} do not expand to the dense planes")
            for kk in wl.owned:
                wl.records(leg, kk)        # raises if a count exceeded the staging capacity
    valid_fraction = float((ctx.download(wl.owned[len(wl.owned) // 2])[0] > 1e-6).mean())

    # per-kernel durations: one more pass of the same step with event timing on, then one with every stream drained
    # before each kernel (stand-alone durations)
    ctx.timing_enable(True)
    ctx.timing_reset()
    wl.step("value")
    ctx.sync()
    kt = ctx.timing()
    ctx.timing_enable(True, serial=True)
    ctx.timing_reset()
    wl.step("value")
    ctx.sync()
    kt_alone = ctx.timing()
    ctx.timing_enable(False)

    kf_total_per_step = args.nkf * world
    rate = {leg: kf_total_per_step * args.steps / (ms[leg] * 1e-3) for leg in LEGS}

    # one more pass of the value step with CUDA events on the compute stream (fusion / intra stage / NCCL) after every
    # op: when the boundary key frames' maps are ready, when the exchange has completed, when the last map is ready
    timeline = None
    if world > 1:
        st = torch.cuda.ExternalStream(ctx.compute_stream(), device=dev)
        marks = []

        def mark(label):
            e = torch.cuda.Event(enable_timing=True)
            e.record(st)
            marks.append((label, e, time.perf_counter()))
        ctx.sync()
        barrier()
        t_host0 = time.perf_counter()
        mark("start")
        wl.step("value", mark=mark)
        ctx.sync()
        labels = ["start", "uploads_issued", "boundary_maps_ready", "exchange_done", "last_map_ready"]
        timeline = {"stream": "compute stream (k1_fuse, k23_intra, NCCL send/recv); K4 and the search run on their own streams",
                    "device_ms": {lab: round(marks[0][1].elapsed_time(m[1]), 3) for lab, m in zip(labels, marks)},
                    "host_issue_ms": {lab: round(1e3 * (m[2] - t_host0), 3) for lab, m in zip(labels, marks)}}

    mine = {"rank": rank, "timeline": timeline, "ms_per_step": {k: round(v, 3) for k, v in rank_ms.items()},
            "host_issue_ms_per_step": {k: round(v, 3) for k, v in host_issue_ms.items()}, "clocks": rank_clocks, "numa_cpus": numa}
    per_rank = [mine]
    if world > 1:
        per_rank = [None] * world
        dist.all_gather_object(per_rank, mine)

    seam = None
    if world > 1 and not args.no_seam_check:
        seam = seam_parity(args, torch, dist, dev, pkg, synth, wl, world, rank)

    line = None
    if rank == 0:
        peak, peak_src, sm_max_mhz = measured_peaks()
        px = W * H
        bpp = algorithmic_bytes_per_px(n)
        k1_ms, k1_n = kt["search"]          # stage K1 = k1_search + k1_fuse (one launch each per key frame)
        k1_ms += kt["fuse"][0]
        k1_avg_s = (k1_ms / max(k1_n, 1)) * 1e-3
        achieved = bpp["search"] * px / k1_avg_s / 1e9 if k1_avg_s > 0 else 0.0
        alone_us = {k: (v[0] / max(v[1], 1)) * 1e3 for k, v in kt_alone.items()}
        k1_alone_s = (alone_us["search"] + alone_us["fuse"]) * 1e-6
        prof = {}
        try:
            with open(os.path.join(ROOT, "profiles", "k1_traffic.json")) as f:
                prof = json.load(f)
        except Exception:
            pass
        if (prof.get("camera", "720p"), int(prof.get("neighbours", 10))) != (args.camera, n):
            prof = {}          # the ncu counters were taken on another workload: no traffic / issue figure for this one
        value = rate["value"]
        sm_mhz = (clocks or {}).get("sm_mhz") or sm_max_mhz
        issue = None
        if prof.get("warp_instructions_per_keyframe"):
            # the bound that binds: warp instructions the path issues per key frame against the machine's
            # issue slots (148 SMs x 4 schedulers x SM clock) over the time a key frame takes in the timed region
            wi = float(prof["warp_instructions_per_keyframe"])
            slots = 148 * 4 * sm_mhz * 1e6 * (1.0 / (value / world))
            issue = {"bound": "issue", "warp_instructions_per_keyframe": wi, "issue_slots_per_keyframe": slots,
                     "frac": wi / slots, "sm_mhz": sm_mhz, "source": prof.get("warp_instructions_source"),
                     "note": "instruction count = AVERAGE over the key frames of the ncu launch list (k1_search 155 M per launch); the "
                             "single --set full capture of one k1_search launch (profiles/*_k1_search_ncu_full.json, 134.7 M) is one "
                             "key frame early in the sequence and is not what this fraction uses"}

        def e2e_entry(leg):
            h2d, d2h = bytes_ps[leg]
            return {"value": rate[leg], "unit": UNIT, "ms_per_step": ms[leg] / args.steps, "h2d_bytes_per_step": h2d,
                    "d2h_bytes_per_step": d2h, "host_issue_ms_per_step": host_issue_ms[leg]}

        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms["value"] / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": config_block(args, world, W, H),
            "e2e_result": "compact records {pixel, inverse depth, sigma} of every pixel that carries a depth or a sigma (both "
                          "reference maps, PM.cc:225-226; the consumer expands them) to pinned host memory; e2e_points adds "
                          "the world point of every record (PM.cc:361-363); e2e_dense = the two maps as dense planes",
            "pixels_per_s": value * px,
            "e2e": e2e_entry("e2e"), "e2e_points": e2e_entry("e2e_points"), "e2e_dense": e2e_entry("e2e_dense"),
            "gpu_launches": int(launches["value"]),
            "launches_per_keyframe": launches["value"] / (args.steps * len(wl.owned)),
            "host_issue_ms_per_step": host_issue_ms,
            "clocks": clocks,
            "roofline": {"kernel": "k1_search + k1_fuse (epipolar search + hypothesis fusion)", "bound": "hbm", "achieved": achieved,
                         "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": prof.get("dram_bytes_per_launch"), "traffic_source": prof.get("source"),
                         "traffic_measured_at_commit": prof.get("commit"),
                         "algorithmic_bytes_per_launch": bpp["search"] * px,
                         "avg_launch_us": k1_avg_s * 1e6,
                         "standalone_launch_us": k1_alone_s * 1e6,
                         "standalone_achieved": bpp["search"] * px / k1_alone_s / 1e9 if k1_alone_s > 0 else 0.0,
                         "whole_path_frac": sum(bpp.values()) * px * (value / world) / 1e9 / peak,
                         "issue": issue},
            "kernels_avg_us": {k: (v[0] / max(v[1], 1)) * 1e3 for k, v in kt.items()},
            "kernels_standalone_us": alone_us,
            "kernels_standalone_share": {k: v / max(sum(alone_us.values()), 1e-9) for k, v in alone_us.items()},
            "valid_fraction": valid_fraction,
            # every rank's own span (the headline figures are the max over ranks), its host issue time and the clocks /
            # power / throttle reasons of ITS GPU during the value and e2e legs: tells GPU-to-GPU variance from a schedule effect
            "per_rank": per_rank,
        }
        if seam is not None:
            line["seam_parity"] = seam
    class_scene = None
    if rank == 0 and world == 1 and not args.no_class_leg:
        import tempfile
        class_scene = os.path.join(tempfile.mkdtemp(prefix="sdpm_class_"), "scene.bin")
        write_class_scene(class_scene, synth, wl, n)
    wl = None
    ctx.close()
    if class_scene:
        line["e2e_class"] = class_leg(args, class_scene, args.nkf)
    if rank == 0 and world == 1 and not args.no_configs:
        line["configs"] = other_configs(args, torch, pkg, synth, sharding, dev)
    if world > 1 and not args.no_configs:
        cfgs = sharded_configs(args, torch, dist, dev, pkg, synth, sharding, world, rank, local)
        if rank == 0:
            line["configs"] = cfgs
    if rank == 0 and world > 1 and not args.no_cpu_baseline:   # every line carries the CPU leg; the other ranks wait at the barrier below
        from oracle import pm_oracle as po
        threads = host_cores()
        sample = max(args.cpu_sample, n + 1)
        _, plan, images = cpu_sample_plan(args, synth, po, sample)
        dt = cpu_pass(po, plan, images, threads)
        line["cpu_baseline"] = {"value": sample / dt, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"first {sample} key frames of the same workload, all stages K0..K5, oracle port with OpenMP on the reference's five pragma loops (one pass, {dt:.1f} s), rank 0's host only"}
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import pm_oracle as po
        threads = host_cores()
        sample = max(args.cpu_sample, n + 1)
        _, plan, images = cpu_sample_plan(args, synth, po, sample)
        dt = cpu_pass(po, plan, images, threads)
        line["cpu_baseline"] = {"value": sample / dt, "unit": UNIT, "cores": threads, "kind": "port",
                                "sample": f"first {sample} key frames of the same workload, all stages K0..K5, oracle port with OpenMP on the reference's five pragma loops (one pass, {dt:.1f} s)"}
        # "as shipped": the reference builds without -fopenmp -> one thread, on the smallest sample that still
        # has n neighbours per key frame; and BASELINE config 1 in full (the reference's own CPU-runnable case), both ways
        s1 = n + 1
        _, plan1, images1 = cpu_sample_plan(args, synth, po, s1)
        dt1 = cpu_pass(po, plan1, images1, 1)
        line["cpu_baseline_1thread"] = {"value": s1 / dt1, "unit": UNIT, "cores": 1, "kind": "port",
                                        "sample": f"first {s1} key frames of the same workload, one pass ({dt1:.1f} s)"}
        if "configs" in line and "config1_480p_n6" in line["configs"]:
            a1 = argparse.Namespace(**dict(vars(args), camera="480p", nbr=6, prior_sigmas=2.0))
            _, planc, imagesc = cpu_sample_plan(a1, synth, po, 10)
            dtc, dtc1 = cpu_pass(po, planc, imagesc, threads), cpu_pass(po, planc, imagesc, 1)
            line["configs"]["config1_480p_n6"]["cpu_baseline"] = {"value": 10 / dtc, "cores": threads, "value_1thread": 10 / dtc1,
                                                                  "unit": UNIT, "kind": "port", "sample": "all 10 key frames, one pass each"}
    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def seam_parity(args, torch, dist, dev, pkg, synth, wl, world, rank, only=None, threads=None):
    """Every rank: the owned key frames with neighbours on another rank -- the first and last N/2 of its range -- against the
    CPU oracle run on the window around the seam, DEPTH / SIGMA / DEPTH_CHECKED bitwise (the loop being sharded is
    PM.cc:262-307, the reads across the seam PM.cc:705-752).  Returns the all-reduced counts."""
    from oracle import pm_oracle as po
    ctx, n, total = wl.ctx, wl.n, wl.total_kf
    own = set(wl.owned)
    seam = [k for k in wl.owned if any(int(j) not in own for j in wl.plan.nbrs[k])]
    if only is not None:   # a subset (long neighbour lists make the oracle window expensive): only(rank, seam) -> key frames
        seam = list(only(rank, seam))
    plan = HostPlan(synth, wl.scene, total, n, [], po.stereo_search_constraints, args.prior_sigmas)
    need_sd = set(seam)
    for k in seam:
        need_sd.update(int(j) for j in wl.plan.nbrs[k])
    for k in sorted(need_sd):
        plan.add(synth, total, n, k, po.stereo_search_constraints, args.prior_sigmas)
    need_img = set(need_sd)
    for k in need_sd:
        need_img.update(int(j) for j in plan.nbrs[k])
    hnp = wl.host_img.numpy()
    images = {k: (hnp[wl.slot[k]] if k in wl.slot else wl.scene.render(k)) for k in sorted(need_img)}
    ora = po.Oracle(threads=threads or max(1, host_cores() // world))
    for k in sorted(need_img):
        ora.add_keyframe(k, images[k], wl.scene.K, wl.scene.pose(k))
    for k in sorted(need_sd):
        ora.semidense(k, plan.nbrs[k], plan.rots[k], *plan.prior[k])
    bad = 0
    for k in seam:
        ora.interkf(k, plan.nbrs[k])
        for which in (po.DEPTH, po.SIGMA, po.DEPTH_CHECKED):
            if not np.array_equal(ctx.plane(k, which), ora.plane(k, which)):
                bad += 1
                sys.stderr.write(f"rank {rank}: seam key frame {k} plane {which} differs from the oracle\n")
    ora.close()
    t = torch.tensor([len(seam), bad], dtype=torch.int64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return {"keyframes_checked": int(t[0].item()), "planes": 3, "mismatching": int(t[1].item()),
            "checker": "CPU oracle on the +-N window around every seam, bitwise", "ranks": world}


def sharded_configs(args, torch, dist, dev, pkg, synth, sharding, world, rank, local):
    """world > 1: BASELINE configs[2] AS STATED -- 1 000 key frames in total, N = 20, sharded over the ranks (strong
    scaling: 1000 / world key frames per rank, a 10-key-frame halo per side) -- and configs[4] (1080p, prior widened to
    mean +- 4 sigma, N = 5 and N = 40, 24 key frames per rank), 2 timed steps each after 1 warm-up, device time = max over
    ranks of the span over all of the context's streams.  Config 3 carries its own seam check: the key frame on either side
    of the seam between ranks 0 and 1 against the CPU oracle (a 21-key-frame window each at N = 20)."""
    peak, _, _ = measured_peaks()
    out = {}
    for name, cam, per_rank, n, sig, scaling in (("config3_720p_n20_1000kf", "720p", 1000 // world, 20, 2.0, "strong"),
                                                 ("config5_1080p_n5", "1080p", 24, 5, 4.0, "weak"),
                                                 ("config5_1080p_n40", "1080p", 24, 40, 4.0, "weak")):
        c = pkg.Context(local)
        ids = [pkg.nccl_unique_id() if rank == 0 else None]
        dist.broadcast_object_list(ids, src=0, device=dev)
        c.comm_init(world, rank, ids[0])
        w = Workload(torch, pkg, synth, sharding, c, dev, cam, per_rank, n, world, rank, args.seed, sig, legs=("value", "e2e"), record_fill=0.6)
        res, steps = {}, 2
        for leg in ("value", "e2e"):
            w.step(leg)
            c.sync()
            dist.barrier()
            torch.cuda.synchronize()
            c.span_begin()
            for _ in range(steps):
                w.step(leg)
            t = torch.tensor([c.span_end()], dtype=torch.float64, device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            res[leg] = per_rank * world * steps / (float(t.item()) * 1e-3)
        for kk in w.owned:
            w.records("e2e", kk)           # raises if a count exceeded the staging capacity
        bpp = sum(algorithmic_bytes_per_px(n).values())
        halo = sorted({k for ids_ in w.shard.recv.values() for k in ids_})
        out[name] = {"camera": cam, "keyframes_total": per_rank * world, "keyframes_per_rank": per_rank, "halo_keyframes_rank0": len(halo),
                     "neighbours": n, "prior_sigmas": sig, "steps": steps, "warmup": 1, "scaling": scaling, "n_gpus": world,
                     "value": res["value"], "e2e": res["e2e"], "unit": UNIT,
                     "whole_path_frac_per_gpu": bpp * w.W * w.H * res["value"] / world / 1e9 / peak}
        if name.startswith("config3") and not args.no_seam_check:
            a3 = argparse.Namespace(**dict(vars(args), prior_sigmas=sig))
            pick = lambda r, seam: ([seam[-1]] if r == 0 else [seam[0]]) if r < 2 and seam else []
            out[name]["seam_parity"] = seam_parity(a3, torch, dist, dev, pkg, synth, w, world, rank, only=pick,
                                                   threads=max(1, host_cores() // 2))
        w = None
        c.comm_destroy()
        c.close()
        dist.barrier()
    return out


def write_class_scene(path, synth, wl, n):
    """The bench workload as a scene file of host/test_pm_class.cc (format in its header): same images, poses, neighbour
    lists and sparse depths as the C-ABI legs; key points such that GetRotInPlane's median is the trajectory's roll."""
    import struct
    scene, ids = wl.scene, list(wl.owned)
    index = {k: i for i, k in enumerate(ids)}
    hnp = wl.host_img.numpy()
    with open(path, "wb") as f:
        f.write(struct.pack("<4i", len(ids), scene.w, scene.h, n))
        for k in ids:
            f.write(struct.pack("<i", k))
            f.write(np.asarray(scene.K, np.float32).tobytes())
            f.write(np.asarray(wl.poses[k], np.float32).tobytes())
            cov = [index[int(j)] for j in wl.plan.nbrs[k]]
            f.write(struct.pack("<i", len(cov)))
            f.write(np.asarray(cov, np.int32).tobytes())
            d = np.asarray(scene.sparse_depths(k, 1000), np.float32)
            f.write(struct.pack("<i", len(d)))
            f.write(d.tobytes())
            mp = np.arange(50, dtype=np.int64)
            ang = np.full(50, 100.0, np.float32) + np.float32(synth.Scene.rot_in_plane(0, k))
            f.write(struct.pack("<i", len(mp)))
            f.write(mp.tobytes())
            f.write(ang.tobytes())
            f.write(hnp[wl.slot[k]].tobytes())
        f.write(struct.pack("<i", 0))


def class_leg(args, scene_path, nkf):
    """e2e_class: the same workload through the drop-in C++ class (host/ProbabilityMapping.cc), timed by its own driver
    with the host clock around whole passes: upload of every image, SemiDenseLoop in windows over sd_semidense_loop,
    compact records with points downloaded asynchronously and EXPANDED into the dense depth_map_ / depth_sigma_ /
    SemiDensePointSets_ members of every key frame (what PM.cc:225-226, 361-363 leave on the host)."""
    import subprocess
    import tempfile
    driver = os.path.join(ROOT, "orb-slam-free-space-carving_b200", "lib", "test_pm_class")
    out_path = os.path.join(os.path.dirname(scene_path), "class_out.bin")
    steps = max(2, min(args.steps, 5))
    try:
        r = subprocess.run([driver, f"--bench={steps},2,{args.class_window},{args.class_threads}", scene_path, out_path], capture_output=True, text=True, timeout=600)
        got = [ln for ln in r.stdout.splitlines() if ln.startswith("bench: ")]
        if r.returncode != 0 or not got:
            return {"unavailable": (r.stderr or r.stdout).strip()[-300:]}
        b = json.loads(got[0][len("bench: "):])
        sys.stderr.write("".join(ln + "\n" for ln in r.stdout.splitlines() if ln.startswith("last pass")))
    finally:
        for q in (scene_path, out_path):
            if os.path.exists(q):
                os.remove(q)
    return {"value": b["kf_per_s"], "unit": UNIT, "ms_per_step": b["ms_per_step"], "steps": b["steps"], "warmup": b["warmup"],
            "window": b["window"], "inter_checked": b["inter_checked"], "keyframes": nkf, "clock": "host steady_clock around whole passes",
            "what": "ProbabilityMapping::SemiDenseLoop (C++ drop-in class over the C ABI): image upload, windows of sd_semidense_loop, "
                    "asynchronous compact downloads with points, records expanded into the dense KeyFrame members of EVERY key frame "
                    "by host threads while the GPU runs the next window"}


def other_configs(args, torch, pkg, synth, sharding, dev):
    """Short runs (2 steps) of the other BASELINE.json configurations on one GPU, so that their numbers are the driver's:
    configs[0] (10 x 640x480, N=6), configs[2] shape (720p, N=20), configs[4] shapes (1080p, N=5 and N=40, prior widened
    to mean +- 4 sigma)."""
    peak, _, _ = measured_peaks()
    out = {}
    for name, cam, nkf, n, sig in (("config1_480p_n6", "480p", 10, 6, 2.0), ("config3_720p_n20", "720p", 48, 20, 2.0),
                                   ("config5_1080p_n5", "1080p", 40, 5, 4.0), ("config5_1080p_n40", "1080p", 48, 40, 4.0)):
        c = pkg.Context(dev.index or 0)
        w = Workload(torch, pkg, synth, sharding, c, dev, cam, nkf, n, 1, 0, args.seed, sig, legs=("value", "e2e"), record_fill=0.6)
        res = {}
        for leg in ("value", "e2e"):
            for _ in range(2):
                w.step(leg)
                c.sync()
            c.span_begin()
            steps = 2 if nkf >= 40 else 20
            for _ in range(steps):
                w.step(leg)
            res[leg] = nkf * steps / (c.span_end() * 1e-3)
        bpp = sum(algorithmic_bytes_per_px(n).values())
        out[name] = {"camera": cam, "keyframes": nkf, "neighbours": n, "prior_sigmas": sig, "steps": steps,
                     "value": res["value"], "e2e": res["e2e"], "unit": UNIT, "pixels_per_s": res["value"] * w.W * w.H,
                     "whole_path_frac": bpp * w.W * w.H * res["value"] / 1e9 / peak,
                     "d2h_bytes_per_step": w.bytes_per_step("e2e")[1]}
        w = None
        c.close()
    return out


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
