"""Developer tool (GPU box): host time to ISSUE the calls of one key frame when the launch queues are empty
(20 key frames after a sync, far below any queue limit) against the device time they take."""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from __graft_entry__ import load_package, PKG_NAME  # noqa: E402
pkg = load_package()
pipeline = importlib.import_module(PKG_NAME + ".pipeline")
synth = importlib.import_module(PKG_NAME + ".synth")
scene = synth.Scene("720p", seed=1234)
nkf, n = 40, 10
plan = pipeline.SequencePlan(scene, nkf, n)
images = {k: scene.render(k) for k in plan.ids}
ctx = pkg.Context(0)
for rep in range(3):
    pipeline.upload_all(ctx, plan, images); ctx.sync()
    t0 = time.perf_counter()
    for k in plan.ids[:20]:
        ctx.semidense(k, plan.nbrs[k], plan.rots[k], *plan.prior[k])
    t1 = time.perf_counter()
    ctx.sync()
    t2 = time.perf_counter()
    for k in plan.ids[20:]:
        ctx.semidense(k, plan.nbrs[k], plan.rots[k], *plan.prior[k])
    ctx.sync()
    t3 = time.perf_counter()
    for k in plan.ids[:20]:
        ctx.interkf_check(k, plan.nbrs[k]); ctx.update_pointset(k)
    t4 = time.perf_counter()
    ctx.sync()
    t5 = time.perf_counter()
    print(f"rep{rep}: semidense issue {1e6*(t1-t0)/20:.1f} us/KF (device done after {1e6*(t2-t0)/20:.1f} us/KF); "
          f"interkf+points issue {1e6*(t4-t3)/20:.1f} us/KF (device {1e6*(t5-t3)/20:.1f})", flush=True)
t0 = time.perf_counter()
for k in plan.ids[:20]:
    ctx.upload_keyframe(k, images[k], scene.K, plan.poses[k])
t1 = time.perf_counter(); ctx.sync(); t2 = time.perf_counter()
print(f"upload issue {1e6*(t1-t0)/20:.1f} us/KF (device {1e6*(t2-t0)/20:.1f})")
ctx.close()
