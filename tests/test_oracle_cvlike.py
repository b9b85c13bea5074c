"""The oracle's restatement of OpenCV's small-matrix arithmetic (oracle/cvlike.h) replayed against
golden vectors produced by the real cv2 (tests/golden/make_cv_golden.py).  Bit-exact."""
import ctypes as C
import os

import numpy as np
import pytest

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "cv_golden.npz"))
fp = C.POINTER(C.c_float)


def p(a):
    return a.ctypes.data_as(fp)


def c32(a):
    return np.ascontiguousarray(a, np.float32)


def test_fast_atan2(po):
    L = po.lib()
    got = np.array([L.ora_fast_atan2(float(y), float(x)) for y, x in zip(G["atan_y"], G["atan_x"])], np.float32)
    assert np.array_equal(got, G["atan_out"])
    assert got.min() >= 0 and got.max() <= 360


def test_gemm_small_33x31(po):
    L = po.lib()
    for i in range(len(G["g31_A"])):
        A, b, c = c32(G["g31_A"][i]), c32(G["g31_b"][i]), c32(G["g31_c"][i])
        d = np.empty(3, np.float32)
        L.ora_cv_gemm_small_33x31(p(A), p(b), float(G["g31_alpha"][i]), p(c), 1.0, p(d))
        assert np.array_equal(d, G["g31_out"][i][:, 0]), i
        L.ora_cv_gemm_small_33x31(p(A), p(b), -1.0, None, 0.0, p(d))
        assert np.array_equal(d, G["g31_out_noc"][i][:, 0]), i


def test_gemm_small_33x33_and_ABt(po):
    L = po.lib()
    for i in range(len(G["g31_A"])):
        A, B = c32(G["g31_A"][i]), c32(G["g33_B"][i])
        D = np.empty((3, 3), np.float32)
        L.ora_cv_gemm_small_33x33(p(A), p(B), p(D))
        assert np.array_equal(D, G["g33_out"][i]), i
        L.ora_cv_gemm_acc64_ABt(p(A), p(B), 1.0, p(D))
        assert np.array_equal(D, G["gabt_out"][i]), i
        L.ora_cv_gemm_acc64_ABt(p(A), p(B), -1.0, p(D))
        assert np.array_equal(D, G["gabt_out_neg"][i]), i


def test_gemm_acc64_dot(po):
    L = po.lib()
    for i in range(len(G["dot_a"])):
        n = int(G["dot_n"][i])
        a, b = c32(G["dot_a"][i][:n]), c32(G["dot_b"][i][:n])
        got = L.ora_cv_gemm_acc64_dot(p(a), p(b), n, float(G["dot_alpha"][i]))
        assert np.float32(got) == G["dot_out"][i], i


def test_gemm_small_44x41(po):
    L = po.lib()
    for i in range(len(G["g41_A"])):
        A, b = c32(G["g41_A"][i]), c32(G["g41_b"][i])
        d = np.empty(4, np.float32)
        L.ora_cv_gemm_small_44x41(p(A), p(b), p(d))
        assert np.array_equal(d, G["g41_out"][i][:, 0]), i


def test_invert_and_solve(po):
    L = po.lib()
    for i in range(len(G["inv_A"])):
        A = c32(G["inv_A"][i])
        D = np.empty((3, 3), np.float32)
        L.ora_cv_invert33(p(A), p(D))
        assert np.array_equal(D, G["inv_out"][i]), i
        At, B = c32(G["solve_A"][i]), c32(G["g33_B"][i])
        L.ora_cv_solve_lu33(p(At), p(B), p(D))
        assert np.array_equal(D, G["solve_out"][i]), i


def test_front_end_against_cv2(po):
    """GradImg / GradTheta / I_stddev definition against cv2.Sobel, cv2.fastAtan2, cv2.meanStdDev."""
    img = G["fe_img"]
    ora = po.Oracle()
    K = np.array([50, 50, 26, 18], np.float32)
    T = np.eye(4, dtype=np.float32)[:3].reshape(-1)
    ora.add_keyframe(0, img, K, T)
    gx, gy = G["fe_gx"], G["fe_gy"]
    assert np.array_equal(ora.plane(0, po.GRAD), np.sqrt(gx * gx + gy * gy).astype(np.float32))
    L = po.lib()
    th = np.array([[L.ora_fast_atan2(float(gy[r, c]), float(gx[r, c])) for c in range(img.shape[1])] for r in range(img.shape[0])], np.float32)
    assert np.array_equal(ora.plane(0, po.THETA), th)
    assert ora.istddev(0) == np.float32(G["fe_std"])
