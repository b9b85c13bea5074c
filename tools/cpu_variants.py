"""Developer tool (GPU box): the CPU baseline variants of BASELINE.md section 3 -- the oracle port with 1 thread
("as shipped": the reference never passes -fopenmp) and with all host threads ("as intended") -- on config 1 in full
and on a 12-key-frame slice of config 2."""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package, PKG_NAME  # noqa: E402
load_package()
pipeline = importlib.import_module(PKG_NAME + ".pipeline")
synth = importlib.import_module(PKG_NAME + ".synth")
from oracle import pm_oracle as po  # noqa: E402

for cam, nkf, n in (("480p", 10, 6), ("720p", 12, 10)):
    scene = synth.Scene(cam, seed=1234)
    plan = pipeline.SequencePlan(scene, nkf, n)
    images = {k: scene.render(k) for k in plan.ids}
    for threads in (po.lib().ora_max_threads(), 1):
        ora = po.Oracle(threads=threads)
        t0 = time.perf_counter()
        for k in plan.ids:
            ora.add_keyframe(k, images[k], scene.K, plan.poses[k])
        for k in plan.ids:
            ora.semidense(k, plan.nbrs[k], plan.rots[k], *plan.prior[k])
        for k in plan.ids:
            ora.interkf(k, plan.nbrs[k])
            ora.pointset(k)
        dt = time.perf_counter() - t0
        ora.close()
        print(f"{cam} {nkf} KF N={n}: {threads:2d} thread(s) {dt:7.2f} s -> {nkf / dt:.3f} KF/s, {nkf * scene.w * scene.h / dt / 1e6:.2f} Mpx/s", flush=True)
