"""K6, the confident-point feed: sd_confident_points against the oracle's restatement of
the selection of PM.cc:114-127.  Index decisions (which pixels, in which order, which survive the smallest-sigma
cut) must be bit-exact; the points follow the point-set tolerance (measured: bit-identical)."""
import numpy as np
import pytest

from helpers import SMALL_CAM, run_gpu, run_oracle

pytestmark = pytest.mark.gpu


def _same(g, o, what):
    gx, gs, gp, gn = g
    ox, os_, op, on = o
    assert gn == on, f"{what}: count {gn} vs {on}"
    assert np.array_equal(gp, op), f"{what}: selected pixels differ"
    assert np.array_equal(gs.view(np.uint32), os_.view(np.uint32)), f"{what}: sigma differs"
    assert np.array_equal(gx.view(np.uint32), ox.view(np.uint32)), f"{what}: points differ"


def test_confident_points_sequence(pkg, po, synth, pipeline):
    scene = synth.Scene(SMALL_CAM, seed=21)
    plan = pipeline.SequencePlan(scene, 8, 6)
    images = {k: scene.render(k) for k in plan.ids}
    ora = run_oracle(po, plan, images)
    ctx = run_gpu(pkg, pipeline, plan, images)
    for k in (0, 3, 7):
        sig = ora.plane(k, po.SIGMA)
        valid = ora.plane(k, po.DEPTH_CHECKED) > 1e-6
        assert valid.sum() > 1000
        med = float(np.median(sig[valid]))
        for max_sigma in (0.01, med, 1e9, 0.0):
            full = ora.confident_points(k, max_sigma, 0)
            _same(ctx.confident_points(k, max_sigma, 0), full, f"kf {k} max_sigma {max_sigma} all")
            n = full[3]
            for max_points in (1, 7, n // 3, n - 1, n, n + 5):
                if max_points <= 0:
                    continue
                _same(ctx.confident_points(k, max_sigma, max_points), ora.confident_points(k, max_sigma, max_points),
                      f"kf {k} max_sigma {max_sigma} max_points {max_points}")
    # the list equals a host-side filter of the downloaded planes (what SavePointCloudObj used to do)
    d, s, p = ctx.download(3, points=True)
    keep = ~(s.astype(np.float64) > 0.01) & (d > 1e-6)
    xyz, sg, pix, n = ctx.confident_points(3, 0.01, 0)
    assert n == int(keep.sum())
    assert np.array_equal(xyz, p.reshape(scene.h, scene.w, 3)[keep])
    # capacity smaller than the selection: count still reports the selection
    xyz2, _, _, n2 = ctx.confident_points(3, 1e9, 0, capacity=10)
    assert n2 == ora.confident_points(3, 1e9, 0)[3] and len(xyz2) == 10
    # pose refresh (PM.cc:321-334): the feed follows sd_set_pose without re-running the mapping
    T = np.array(plan.poses[3], np.float32).reshape(3, 4).copy()
    T[:, 3] += np.array([0.01, -0.02, 0.03], np.float32)
    ctx.set_pose(3, T)
    ora.set_pose(3, T)
    _same(ctx.confident_points(3, 0.01, 500), ora.confident_points(3, 0.01, 500), "after set_pose")
    # device-resident variant: same counts, pointers set
    px, ps, pp, n3, nconf = ctx.confident_points_device(3, 0.01, 500)
    assert px and ps and pp and n3 == min(500, nconf) and nconf == ora.confident_points(3, 0.01, 0)[3]
    ctx.close()


@pytest.mark.parametrize("w,h", [(320, 240), (517, 61), (1283, 17)])
def test_confident_points_ties_and_odd_values(pkg, po, w, h):
    """Crafted maps: heavily tied sigmas (the cut falls inside a run of equal keys), negative, zero, inf and NaN
    sigmas, depth on the image border, widths that are not a multiple of the 256-pixel segment."""
    import torch
    rng = np.random.default_rng(w + h)
    img = rng.integers(0, 256, (h, w)).astype(np.uint8)
    K = np.array([300, 310, w / 2 - 0.5, h / 2 - 0.5], np.float32)
    c, s_ = np.cos(0.1), np.sin(0.1)
    T = np.array([[c, -s_, 0, 0.1], [s_, c, 0, -0.2], [0, 0, 1, 0.05]], np.float32)
    depth = np.where(rng.random((h, w)) < 0.4, rng.uniform(0.2, 2.0, (h, w)), 0).astype(np.float32)
    sigma = (rng.integers(0, 12, (h, w)) * 0.0025).astype(np.float32)          # 12 distinct values, many ties
    odd = rng.random((h, w))
    sigma[odd < 0.02] = np.nan
    sigma[(odd >= 0.02) & (odd < 0.04)] = -0.005
    sigma[(odd >= 0.04) & (odd < 0.05)] = np.inf
    sigma[(odd >= 0.05) & (odd < 0.06)] = -np.nan
    ctx = pkg.Context(0)
    ora = po.Oracle()
    ctx.upload_keyframe(0, img, K, T.reshape(-1))
    ora.add_keyframe(0, img, K, T.reshape(-1))
    dd = torch.from_numpy(depth).cuda()
    ss = torch.from_numpy(sigma).cuda()
    ctx.import_maps(0, dd.data_ptr(), ss.data_ptr())
    ctx.sync()
    ora.set_maps(0, depth, sigma)
    # K5 on the same crafted map: widths that are not a multiple of 4 take the scalar store path of k5_pointset
    ctx.update_pointset(0)
    ora.pointset(0)
    assert np.array_equal(ctx.plane(0, po.POINTS).view(np.uint32), ora.plane(0, po.POINTS).view(np.uint32))
    for max_sigma in (0.01, 0.0125, 1e30):
        full = ora.confident_points(0, max_sigma, 0)
        _same(ctx.confident_points(0, max_sigma, 0), full, f"max_sigma {max_sigma} all")
        n = full[3]
        for max_points in sorted({1, 2, 33, n // 7, n // 2, n - 3, n - 1}):
            if 0 < max_points:
                _same(ctx.confident_points(0, max_sigma, max_points), ora.confident_points(0, max_sigma, max_points),
                      f"max_sigma {max_sigma} max_points {max_points}")
    ctx.close()
