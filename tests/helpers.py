"""Shared by the parity tests: run the oracle / the CUDA path over a SequencePlan and compare."""
import numpy as np

SMALL_CAM = dict(w=320, h=240, fx=262.5, fy=262.5, cx=159.5, cy=119.5)


def run_oracle(po, plan, images, threads=None, params=None, pointset=True, interkf=True):
    ora = po.Oracle(params=params, threads=threads or po.lib().ora_max_threads())
    for k in plan.ids:
        ora.add_keyframe(k, images[k], plan.scene.K, plan.poses[k])
    for k in plan.ids:
        ora.semidense(k, plan.nbrs[k], plan.rots[k], *plan.prior[k])
    if interkf:
        for k in plan.ids:
            ora.interkf(k, plan.nbrs[k])
            if pointset:
                ora.pointset(k)
    return ora


def run_gpu(pkg, pipeline, plan, images, params=None, pointset=True, interkf=True, ctx=None):
    ctx = ctx or pkg.Context(0, params)
    pipeline.upload_all(ctx, plan, images)
    pipeline.semidense_all(ctx, plan)
    if interkf:
        pipeline.interkf_all(ctx, plan, pointset=pointset)
    ctx.sync()
    return ctx


def compare_maps(g, o, what=""):
    """north_star tolerance: valid mask bit-exact up to <= 0.1 % float-boundary flips; values within
    1e-4 relative where both are valid.  Returns (flips, max_rel, n_exact_diff)."""
    mg, mo = g > 1e-6, o > 1e-6
    flips = int(np.count_nonzero(mg != mo))
    both = mg & mo
    rel = float(np.max(np.abs(g[both] - o[both]) / np.abs(o[both]))) if both.any() else 0.0
    assert flips <= 1e-3 * g.size, f"{what}: {flips} valid-mask flips of {g.size}"
    assert rel <= 1e-4, f"{what}: max relative error {rel}"
    return flips, rel, int(np.count_nonzero(g != o))


def assert_sequence_parity(po, ctx, ora, plan, points=True, checked=True):
    stats = {}
    for k in plan.ids:
        assert np.array_equal(ctx.plane(k, po.GRAD), ora.plane(k, po.GRAD)), f"GradImg kf {k}"
        assert np.array_equal(ctx.plane(k, po.THETA), ora.plane(k, po.THETA)), f"GradTheta kf {k}"
        assert ctx.istddev(k) == ora.istddev(k), f"I_stddev kf {k}"
        planes = [po.DEPTH, po.SIGMA] + ([po.DEPTH_CHECKED] if checked else [])
        for which in planes:
            f, r, nd = compare_maps(ctx.plane(k, which), ora.plane(k, which), f"plane {which} kf {k}")
            s = stats.setdefault(which, [0, 0.0, 0])
            s[0] += f
            s[1] = max(s[1], r)
            s[2] += nd
        if points and checked:
            gp, op = ctx.plane(k, po.POINTS), ora.plane(k, po.POINTS)
            gd, od = ctx.plane(k, po.DEPTH_CHECKED), ora.plane(k, po.DEPTH_CHECKED)
            same = np.repeat((gd > 1e-6) & (od > 1e-6), 3, axis=1)
            assert np.allclose(gp[same], op[same], rtol=1e-4, atol=1e-6), f"points kf {k}"
    return stats
