"""Developer tool (GPU box): parity statistics against the oracle + per-kernel timing.
usage: python tools/gpu_check.py [camera] [nkf] [n_nbr] [--no-oracle]"""
import importlib
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package, PKG_NAME  # noqa: E402

pkg = load_package()
pipeline = importlib.import_module(PKG_NAME + ".pipeline")
synth = importlib.import_module(PKG_NAME + ".synth")


def main():
    args = [a for a in sys.argv[1:] if not a.startswith("--")]
    cam = args[0] if len(args) > 0 else "480p"
    nkf = int(args[1]) if len(args) > 1 else 10
    n = int(args[2]) if len(args) > 2 else 6
    use_oracle = "--no-oracle" not in sys.argv
    scene = synth.Scene(cam, seed=1234)
    t = time.time()
    plan = pipeline.SequencePlan(scene, nkf, n)
    images = {k: scene.render(k) for k in plan.ids}
    print(f"[{cam} nkf={nkf} N={n}] render {time.time()-t:.2f}s", flush=True)

    ctx = pkg.Context(0)
    ctx.timing_enable(True)
    for rep in range(3):
        ctx.timing_reset()
        t = time.time()
        pipeline.upload_all(ctx, plan, images)
        pipeline.semidense_all(ctx, plan)
        pipeline.interkf_all(ctx, plan)
        ctx.sync()
        wall = time.time() - t
        tm = ctx.timing()
        tot = sum(v[0] for v in tm.values())
        print(f"rep{rep}: wall {wall*1e3:.1f} ms  kernels {tot:.2f} ms  " +
              "  ".join(f"{k}={v[0]/max(v[1],1)*1e3:.1f}us x{v[1]}" for k, v in tm.items()), flush=True)
    px = scene.w * scene.h
    print(f"KF/s (kernel time) {nkf/(tot*1e-3):.1f}   algorithmic GB/s {(86+17*n)*px*nkf/(tot*1e-3)/1e9:.1f}")
    print("active counts", [ctx.active_count(k) for k in plan.ids[:4]], "of", px)

    if use_oracle:
        from oracle import pm_oracle as po
        ora = po.Oracle(threads=po.lib().ora_max_threads())
        t = time.time()
        for k in plan.ids:
            ora.add_keyframe(k, images[k], scene.K, plan.poses[k])
        for k in plan.ids:
            ora.semidense(k, plan.nbrs[k], plan.rots[k], *plan.prior[k])
        for k in plan.ids:
            ora.interkf(k, plan.nbrs[k])
            ora.pointset(k)
        print(f"oracle ({po.lib().ora_max_threads()} threads) {time.time()-t:.2f}s -> {nkf/(time.time()-t):.2f} KF/s")
        names = {po.GRAD: "grad", po.THETA: "theta", po.DEPTH: "depth", po.SIGMA: "sigma", po.DEPTH_CHECKED: "depth_chk", po.POINTS: "points"}
        for which, nm in names.items():
            nd = nb = 0
            worst = 0.0
            flips = 0
            for k in plan.ids:
                g, o = ctx.plane(k, which), ora.plane(k, which)
                nd += int(np.count_nonzero(g != o))
                nb += g.size
                if which in (po.DEPTH, po.SIGMA, po.DEPTH_CHECKED):
                    mg, mo = g > 1e-6, o > 1e-6
                    flips += int(np.count_nonzero(mg != mo))
                    both = mg & mo
                    if both.any():
                        worst = max(worst, float(np.max(np.abs(g[both] - o[both]) / np.abs(o[both]))))
            print(f"  {nm:10s} differing values {nd} / {nb}   mask flips {flips}   max rel err (both valid) {worst:.3g}")
        print("  istddev equal:", all(ctx.istddev(k) == ora.istddev(k) for k in plan.ids))
        d = ctx.plane(plan.ids[nkf // 2], po.DEPTH_CHECKED)
        print("  valid fraction (mid kf):", float((d > 1e-6).mean()))
    ctx.close()


if __name__ == "__main__":
    main()
