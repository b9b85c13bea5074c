"""Developer tool (GPU box): what each stage costs INSIDE the stream pipeline -- throughput of the step with stages
removed: uploads + K0 only; + search/fuse/intra (no inter-key-frame stage); the full step.
usage: python tools/stage_costs.py [camera] [nkf] [n]"""
import importlib
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from __graft_entry__ import PKG_NAME, load_package  # noqa: E402

pkg = load_package()
synth = importlib.import_module(PKG_NAME + ".synth")
sharding = importlib.import_module(PKG_NAME + ".sharding")


def main():
    cam = sys.argv[1] if len(sys.argv) > 1 else "720p"
    nkf = int(sys.argv[2]) if len(sys.argv) > 2 else 200
    n = int(sys.argv[3]) if len(sys.argv) > 3 else 10
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    ctx = pkg.Context(0)
    wl = bench.Workload(torch, pkg, synth, sharding, ctx, dev, cam, nkf, n, 1, 0, 1234, legs=("value",))
    no_inter = ctx.make_loop([wl._item("value", k) for k in wl.owned], interkf=False)

    def uploads():
        for k in wl.resident:
            ctx.upload_keyframe_ptr(k, wl.dev_img[wl.slot[k]].data_ptr(), wl.W, wl.H, wl.W, wl.scene.K, wl.poses[k], device=True)

    def semi():
        uploads()
        ctx.semidense_loop(no_inter)

    def full():
        wl.step("value")

    import time
    for name, fn in (("K0 only", uploads), ("K0+K1+K23", semi), ("full", full)):
        for _ in range(2):
            fn()
            ctx.sync()
        reps = 5 if nkf >= 100 else 50
        ctx.span_begin()
        for _ in range(reps):
            fn()
        ms = ctx.span_end() / reps
        # host time to issue ONE step into empty queues (no back-pressure from the device)
        ctx.sync()
        t0 = time.perf_counter()
        fn()
        host_us = 1e6 * (time.perf_counter() - t0) / nkf
        ctx.sync()
        print(f"{name:12s} {ms:8.2f} ms/step  {1e3 * ms / nkf:7.1f} us/KF  {nkf / ms * 1e3:8.1f} KF/s   host issue {host_us:6.1f} us/KF", flush=True)
    ctx.close()


if __name__ == "__main__":
    main()
