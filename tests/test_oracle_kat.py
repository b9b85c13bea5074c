"""Known-answer tests of the oracle, derived by hand from the formulas in PM.cc (the reference
ships no tests or vectors for this path), plus ground-truth recovery on a
synthetic scene and a regression golden of the oracle's own output."""
import ctypes as C
import hashlib
import json
import os

import numpy as np
import pytest

from helpers import SMALL_CAM, run_oracle

fp = C.POINTER(C.c_float)
HERE = os.path.dirname(__file__)


def test_fast_atan2_cardinals(po):
    L = po.lib()
    assert L.ora_fast_atan2(0.0, 1.0) == 0.0
    assert abs(L.ora_fast_atan2(1.0, 0.0) - 90.0) < 1e-4
    assert abs(L.ora_fast_atan2(0.0, -1.0) - 180.0) < 1e-4
    assert abs(L.ora_fast_atan2(-1.0, 0.0) - 270.0) < 1e-4
    assert abs(L.ora_fast_atan2(1.0, 1.0) - 45.0) < 0.02   # polynomial accuracy ~0.01 deg
    assert L.ora_fast_atan2(-1e-9, 1.0) == 360.0            # the range is [0, 360], not [0, 360)


def test_bilinear(po):
    """PM.cc:40-59: argument order (img, y, x); integer coordinates return the pixel, half
    coordinates the mean; out-of-range taps clamp to the edge (H2)."""
    L = po.lib()
    img = np.arange(20, dtype=np.uint8).reshape(4, 5) * 10
    p = img.ctypes.data
    assert L.ora_bilinear_u8(p, 5, 4, 2.0, 3.0) == float(img[2, 3])
    assert L.ora_bilinear_u8(p, 5, 4, 1.5, 3.0) == (float(img[1, 3]) + float(img[2, 3])) / 2
    assert L.ora_bilinear_u8(p, 5, 4, 1.0, 2.5) == (float(img[1, 2]) + float(img[1, 3])) / 2
    assert L.ora_bilinear_u8(p, 5, 4, 1.25, 1.0) == 0.75 * img[1, 1] + 0.25 * img[2, 1]
    assert L.ora_bilinear_u8(p, 5, 4, 3.5, 4.0) == float(img[3, 4])       # y1 == rows clamps
    f = img.astype(np.float32)
    assert L.ora_bilinear_f32(f.ctypes.data, 5, 4, 0.5, 0.5) == float(f[:2, :2].mean())


def test_chi_test_threshold(po):
    """PM.cc:920-924: (a-b)^2/sa^2 + (a-b)^2/sb^2 < 5.99"""
    L = po.lib()
    assert L.ora_chi_test(1.0, 1.1, 0.1, 0.1, 5.99) == 1     # 1 + 1 = 2
    assert L.ora_chi_test(1.0, 1.2, 0.1, 0.1, 5.99) == 0     # 4 + 4 = 8
    assert L.ora_chi_test(1.0, 1.0, 0.0, 0.1, 5.99) == 0     # 0/0 -> NaN -> false
    assert L.ora_chi_test(0.5, 0.5, 0.2, 0.3, 5.99) == 1


def _arr(v):
    a = np.asarray(v, np.float32)
    return a, a.ctypes.data_as(fp)


def test_get_fusion(po):
    """PM.cc:947-970: inverse-variance weighted mean; sigma = sqrt(1/sum(1/s^2)); also min sigma."""
    L = po.lib()
    d, dp = _arr([1, 2, 3])
    s, sp = _arr([1, 1, 1])
    od, os_, om = C.c_float(), C.c_float(), C.c_float()
    L.ora_get_fusion(dp, sp, 3, C.byref(od), C.byref(os_), C.byref(om))
    assert od.value == 2.0 and os_.value == np.sqrt(np.float32(1) / np.float32(3)) and om.value == 1.0
    d, dp = _arr([1, 3])
    s, sp = _arr([1, 0.5])
    L.ora_get_fusion(dp, sp, 2, C.byref(od), C.byref(os_), C.byref(om))
    assert od.value == np.float32(13.0 / 5.0) and om.value == 0.5
    L.ora_get_fusion_pairs(dp, sp, 2, C.byref(od), C.byref(os_))     # PM.cc:926-945: sigma = min sigma
    assert od.value == np.float32(13.0 / 5.0) and os_.value == 0.5


def test_hypothesis_fusion_rejects_outlier(po):
    """PM.cc:598-626: largest mutually compatible set (first wins on ties), >= lambdaN to fuse."""
    L = po.lib()
    d, dp = _arr([1.00, 1.01, 0.99, 1.02, 2.0])
    s, sp = _arr([0.02, 0.02, 0.02, 0.02, 0.02])
    od, os_ = C.c_float(), C.c_float()
    assert L.ora_hypothesis_fusion(None, dp, sp, 5, C.byref(od), C.byref(os_)) == 1
    assert abs(od.value - 1.005) < 1e-6 and abs(os_.value - 0.01) < 1e-7
    d, dp = _arr([1.0, 2.0, 3.0, 4.0])                                  # nothing compatible but itself
    s, sp = _arr([0.01] * 4)
    assert L.ora_hypothesis_fusion(None, dp, sp, 4, C.byref(od), C.byref(os_)) == 0
    assert od.value == 0 and os_.value == 0
    assert L.ora_hypothesis_fusion(None, dp, sp, 0, C.byref(od), C.byref(os_)) == 0   # empty list


def test_stereo_search_constraints(po):
    """PM.cc:370-383: max_depth = 1/(mean+2s), min_depth = 1/(mean-2s): an INVERSE-depth interval."""
    mn, mx = po.stereo_search_constraints([1, 1, 1, 1])
    assert mn == 1.0 and mx == 1.0
    mn, mx = po.stereo_search_constraints([0.9, 1.1])
    assert abs(mx - 1 / 1.2) < 1e-6 and abs(mn - 1 / 0.8) < 1e-6 and mn > mx


def test_rot_in_plane_median(po):
    """PM.cc:467-484 + 172-177: angle2-angle1 over shared map points, negative angles skipped,
    lower median, 0 without covisibility."""
    L = po.lib()
    mp1 = np.array([5, -1, 7, 9, 11], np.int64)
    a1, a1p = _arr([10, 20, 30, -1, 50])
    mp2 = np.array([7, 5, 11, 9], np.int64)
    a2, a2p = _arr([33, 14, 58, 40])
    # pairs: 5:(14-10)=4, 7:(33-30)=3, 9: angle1<0 skipped, 11:(58-50)=8 -> sorted [3,4,8] -> 4
    assert L.ora_rot_in_plane_median(mp1.ctypes.data, a1p, 5, mp2.ctypes.data, a2p, 4) == 4.0
    mp3 = np.array([1, 2], np.int64)
    assert L.ora_rot_in_plane_median(mp1.ctypes.data, a1p, 5, mp3.ctypes.data, a2p, 2) == 0.0


def _T(R=np.eye(3), t=(0, 0, 0)):
    return np.hstack([np.asarray(R, np.float32), np.asarray(t, np.float32).reshape(3, 1)]).reshape(-1)


K480 = np.array([525, 525, 319.5, 239.5], np.float32)


def test_pixel_depth_pure_x_translation(po):
    """PM.cc:845-875 with R21 = I, t21 = (tx,0,0) reduces to disparity / (fx * baseline)."""
    L = po.lib()
    F, R, t = po.pair_geometry(K480, _T(), K480, _T(t=(-0.1, 0, 0)))   # camera 2 sits 0.1 to the right
    assert np.array_equal(R, np.eye(3, dtype=np.float32)) and t[0] == np.float32(-0.1)
    _, Rp = _arr(R.reshape(-1))
    _, tp = _arr(t)
    _, Kp = _arr(K480)
    for px, u in ((400, 350.0), (100, 47.5), (320, 300.25)):
        rho = L.ora_pixel_depth(Kp, Rp, tp, u, px, 200)
        assert abs(rho - (u - px) / (525 * -0.1)) < 1e-5 * abs(rho)


def test_search_range_pure_x_translation(po):
    """PM.cc:877-910 incl. the reference's multiplication by INVERSE depth:
    u = fx*(xp*rho + tx)/rho + cx = px + fx*tx/rho."""
    L = po.lib()
    F, R, t = po.pair_geometry(K480, _T(), K480, _T(t=(-0.1, 0, 0)))
    _, Rp = _arr(R.reshape(-1))
    _, tp = _arr(t)
    _, Kp = _arr(K480)
    umin, umax = C.c_float(), C.c_float()
    L.ora_search_range(Kp, Rp, tp, 640, 400, 200, 1.25, 0.8, C.byref(umin), C.byref(umax))
    assert abs(umin.value - (400 - 52.5 / 0.8)) < 1e-3 and abs(umax.value - (400 - 52.5 / 1.25)) < 1e-3
    L.ora_search_range(Kp, Rp, tp, 640, 30, 200, 1.25, 0.8, C.byref(umin), C.byref(umax))
    assert umin.value == 0.0 and umax.value == 0.0                       # clamped to [0, cols]
    _, tp2 = _arr(-t)
    L.ora_search_range(Kp, Rp, tp2, 640, 620, 200, 1.25, 0.8, C.byref(umin), C.byref(umax))
    assert umin.value == 640.0 and umax.value == 640.0                   # clamp is to cols INCLUSIVE (H2)


def test_fundamental_epipolar_constraint(po, synth):
    """PM.cc:972-987: p1^T F12 p2 = 0 for projections of the same 3-D point."""
    sc = synth.Scene("480p")
    rng = np.random.default_rng(5)
    for k1, k2 in ((0, 1), (3, 7), (9, 4)):
        T1, T2 = sc.pose(k1).reshape(3, 4).astype(np.float64), sc.pose(k2).reshape(3, 4).astype(np.float64)
        F, R21, t21 = po.pair_geometry(sc.K, sc.pose(k1), sc.K, sc.pose(k2))
        Kd = np.array([[525, 0, 319.5], [0, 525, 239.5], [0, 0, 1.0]])
        for _ in range(20):
            Xw = np.array([rng.uniform(-0.3, 0.6), rng.uniform(-0.3, 0.3), rng.uniform(0.8, 1.3), 1.0])
            p1, p2 = Kd @ (T1 @ Xw), Kd @ (T2 @ Xw)
            p1, p2 = p1 / p1[2], p2 / p2[2]
            r = p1 @ F.astype(np.float64) @ p2
            scale = np.linalg.norm(F.astype(np.float64).T @ p1)[None][0]   # distance of p2 to the epipolar line of p1, px
            assert abs(r) / scale < 0.05
        # R21/t21 map camera-1 coordinates to camera-2 coordinates
        Xc1 = T1 @ np.array([0.1, -0.05, 1.1, 1.0])
        Xw = np.linalg.solve(np.vstack([T1, [0, 0, 0, 1]]), np.append(Xc1, 1))
        assert np.allclose(R21 @ Xc1 + t21, T2 @ Xw, atol=1e-5)


@pytest.fixture(scope="module")
def small_run(po, synth, pipeline):
    scene = synth.Scene(SMALL_CAM, seed=7)
    plan = pipeline.SequencePlan(scene, 8, 6)
    rendered = {k: scene.render(k, with_depth=True) for k in plan.ids}
    images = {k: v[0] for k, v in rendered.items()}
    gt = {k: v[1] for k, v in rendered.items()}
    ora = run_oracle(po, plan, images)
    return scene, plan, images, gt, ora


def test_oracle_recovers_ground_truth_depth(po, small_run):
    """Independent pin of the restatement: on a ray-cast scene with known depth the oracle's
    semi-dense maps must be dense enough and metrically right."""
    scene, plan, images, gt, ora = small_run
    for k in plan.ids[2:-2]:
        for which in (po.DEPTH, po.DEPTH_CHECKED):
            d = ora.plane(k, which)
            v = d > 1e-6
            assert v.mean() > 0.05, f"kf {k}: only {v.mean():.3f} valid"
            rel = np.abs(1.0 / d[v] - gt[k][v]) / gt[k][v]
            assert np.median(rel) < 0.005, f"kf {k}: median relative depth error {np.median(rel):.4f}"
            assert np.percentile(rel, 90) < 0.05


def test_oracle_structure_invariants(po, small_run):
    scene, plan, images, gt, ora = small_run
    for k in plan.ids:
        G, d, s, dc = (ora.plane(k, w) for w in (po.GRAD, po.DEPTH, po.SIGMA, po.DEPTH_CHECKED))
        v = d > 1e-6
        assert not v[:2].any() and not v[-2:].any() and not v[:, :2].any() and not v[:, -2:].any()   # 2-px border, PM.cc:198-199
        assert (G[v] >= 8).all()                          # only gated pixels carry depth (growing is a no-op, H7)
        assert (s[v] > 0).all()
        assert not ((dc > 1e-6) & ~v).any()               # the inter-KF check only removes or refines
        P = ora.plane(k, po.POINTS).reshape(scene.h, scene.w, 3)
        assert (P[~(dc > 1e-6)] == 0).all()
        # back-projected points re-project onto their own pixel (PM.cc:352-359)
        T = plan.poses[k].reshape(3, 4).astype(np.float64)
        ys, xs = np.nonzero(dc > 1e-6)
        Xc = (T[:, :3] @ P[ys, xs].astype(np.float64).T).T + T[:, 3]
        u = scene.K[0] * Xc[:, 0] / Xc[:, 2] + scene.K[2]
        vv = scene.K[1] * Xc[:, 1] / Xc[:, 2] + scene.K[3]
        assert np.abs(u - xs).max() < 1e-2 and np.abs(vv - ys).max() < 1e-2


def test_oracle_openmp_is_deterministic(po, small_run):
    """The five `#pragma omp parallel for` loops (PM.cc:197,339,491,554,658) are race free:
    1 thread (as shipped) and many threads give identical planes."""
    scene, plan, images, gt, ora = small_run
    ora1 = run_oracle(po, plan, images, threads=1)
    for k in plan.ids:
        for w in (po.DEPTH, po.SIGMA, po.DEPTH_CHECKED, po.POINTS):
            assert np.array_equal(ora.plane(k, w), ora1.plane(k, w))


def _digest(ora, po, plan):
    h = hashlib.sha256()
    for k in plan.ids:
        for w in (po.GRAD, po.THETA, po.DEPTH, po.SIGMA, po.DEPTH_CHECKED, po.POINTS):
            h.update(ora.plane(k, w).tobytes())
    return h.hexdigest()


def test_oracle_golden_regression(po, small_run):
    """tests/golden/oracle_small.json (written by tests/golden/make_oracle_golden.py) pins the
    oracle's output on the seeded small scene: catches accidental changes of the restatement and
    differences between host CPUs / compilers."""
    scene, plan, images, gt, ora = small_run
    with open(os.path.join(HERE, "golden", "oracle_small.json")) as f:
        g = json.load(f)
    img_digest = hashlib.sha256(b"".join(images[k].tobytes() for k in plan.ids)).hexdigest()
    assert img_digest == g["images_sha256"], "synthetic generator changed: regenerate the golden"
    assert _digest(ora, po, plan) == g["planes_sha256"]
    mid = plan.ids[len(plan.ids) // 2]
    d = ora.plane(mid, po.DEPTH_CHECKED)
    assert int((d > 1e-6).sum()) == g["mid_valid"]
    for (y, x, val) in g["mid_samples"]:
        assert float(d[y, x]) == val
