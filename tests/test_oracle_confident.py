"""Oracle restatement of the confident-point selection (PM.cc:114-127) against a numpy statement of the same rule."""
import numpy as np


def _key(s):
    b = s.view(np.uint32).astype(np.uint64)
    neg = (b & 0x80000000) != 0
    return np.where(neg, (~b) & 0xFFFFFFFF, b | 0x80000000)


def test_confident_points_rule_and_cut(po):
    w, h = 97, 41
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, (h, w)).astype(np.uint8)
    K = np.array([120, 125, 48, 20], np.float32)
    T = np.array([[1, 0, 0, 0.3], [0, 1, 0, -0.1], [0, 0, 1, 0.2]], np.float32)
    depth = np.where(rng.random((h, w)) < 0.5, rng.uniform(0.3, 3.0, (h, w)), 0).astype(np.float32)
    sigma = (rng.integers(0, 8, (h, w)) * 0.004).astype(np.float32)
    sigma[rng.random((h, w)) < 0.03] = np.nan
    ora = po.Oracle()
    ora.add_keyframe(0, img, K, T.reshape(-1))
    ora.set_maps(0, depth, sigma)
    ora.pointset(0)
    P = ora.plane(0, po.POINTS).reshape(h, w, 3)
    keep = ~(sigma.astype(np.float64) > 0.01) & (depth.astype(np.float64) > 0.000001)      # PM.cc:120-121
    xyz, sg, pix, n = ora.confident_points(0, 0.01, 0)
    assert n == int(keep.sum()) and n > 100
    ys, xs = np.nonzero(keep)                                                            # raster order
    assert np.array_equal(pix, (ys.astype(np.uint32) << 16) | xs.astype(np.uint32))
    assert np.array_equal(xyz.view(np.uint32), P[keep].view(np.uint32))
    assert np.array_equal(sg.view(np.uint32), sigma[keep].view(np.uint32))
    # world point of one interior pixel by hand: Twc * (Z(x-cx)/fx, Z(y-cy)/fy, Z, 1), PM.cc:349-363
    i = int(np.flatnonzero((ys > 2) & (ys < h - 3) & (xs > 2) & (xs < w - 3))[0])
    y, x = int(ys[i]), int(xs[i])
    Z = 1.0 / depth[y, x]
    Pc = np.array([Z * (x - K[2]) / K[0], Z * (y - K[3]) / K[1], Z])
    assert np.allclose(xyz[i], T[:, :3].T @ (Pc - T[:, 3]), rtol=1e-5, atol=1e-6)
    # border pixels keep their place in the list but carry the zero point UpdateSemiDensePointSet leaves there
    b = np.flatnonzero((ys < 2) | (xs < 2) | (ys >= h - 2) | (xs >= w - 2))
    assert len(b) and not xyz[b].any()
    # smallest-sigma cut: stable in raster order, result re-emitted in raster order
    for m in (1, 10, n // 2, n - 1):
        order = np.argsort(_key(sigma[keep]), kind="stable")[:m]
        sel = np.sort(order)
        x2, s2, p2, n2 = ora.confident_points(0, 0.01, m)
        assert n2 == m and np.array_equal(p2, pix[sel]) and np.array_equal(x2.view(np.uint32), xyz[sel].view(np.uint32))
    assert ora.confident_points(0, 0.01, n + 10)[3] == n
