"""ctypes binding of the CPU oracle (oracle/pm_oracle.cc).  TEST INFRASTRUCTURE.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module.  The product package never does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "build", "libpm_oracle.so")

GRAD, THETA, DEPTH, SIGMA, DEPTH_CHECKED, POINTS = range(6)


class OraParams(C.Structure):
    _fields_ = [
        ("lambdaG", C.c_float),
        ("lambdaL", C.c_float),
        ("lambdaTheta", C.c_float),
        ("lambdaN", C.c_int),
        ("theta", C.c_double),
        ("chi2_fuse", C.c_double),
        ("chi2_inter", C.c_double),
        ("slope_limit", C.c_float),
        ("border", C.c_int),
        ("interkf_inplace", C.c_int),
    ]


def build(force=False):
    """Compile the oracle with the Makefile in this directory (g++ only)."""
    src = [os.path.join(_HERE, f) for f in ("pm_oracle.cc", "pm_oracle.h", "cvlike.h", "Makefile")]
    if not force and os.path.exists(_LIB_PATH) and all(os.path.getmtime(_LIB_PATH) >= os.path.getmtime(s) for s in src):
        return _LIB_PATH
    subprocess.check_call(["make", "-C", _HERE, "-s", "all"])
    return _LIB_PATH


_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        fp, ip, u8p, vp = C.POINTER(C.c_float), C.POINTER(C.c_int), C.POINTER(C.c_uint8), C.c_void_p
        L.ora_create.restype = vp
        L.ora_create.argtypes = [C.POINTER(OraParams)]
        L.ora_destroy.argtypes = [vp]
        L.ora_default_params.argtypes = [C.POINTER(OraParams)]
        L.ora_set_threads.argtypes = [C.c_int]
        L.ora_max_threads.restype = C.c_int
        L.ora_add_keyframe.argtypes = [vp, C.c_int, vp, C.c_int, C.c_int, C.c_int, fp, fp]
        L.ora_set_pose.argtypes = [vp, C.c_int, fp]
        L.ora_semidense.argtypes = [vp, C.c_int, ip, fp, C.c_int, C.c_float, C.c_float]
        L.ora_interkf.argtypes = [vp, C.c_int, ip, C.c_int]
        L.ora_pointset.argtypes = [vp, C.c_int]
        L.ora_set_maps.argtypes = [vp, C.c_int, vp, vp]
        L.ora_confident_points.argtypes = [vp, C.c_int, C.c_double, C.c_int, C.c_int, vp, vp, vp]
        L.ora_get_plane.argtypes = [vp, C.c_int, C.c_int, vp]
        L.ora_get_istddev.argtypes = [vp, C.c_int]
        L.ora_get_istddev.restype = C.c_float
        L.ora_fast_atan2.argtypes = [C.c_float, C.c_float]
        L.ora_fast_atan2.restype = C.c_float
        L.ora_bilinear_u8.argtypes = [vp, C.c_int, C.c_int, C.c_float, C.c_float]
        L.ora_bilinear_u8.restype = C.c_float
        L.ora_bilinear_f32.argtypes = [vp, C.c_int, C.c_int, C.c_float, C.c_float]
        L.ora_bilinear_f32.restype = C.c_float
        L.ora_chi_test.argtypes = [C.c_float, C.c_float, C.c_float, C.c_float, C.c_double]
        L.ora_get_fusion.argtypes = [fp, fp, C.c_int, fp, fp, fp]
        L.ora_get_fusion_pairs.argtypes = [fp, fp, C.c_int, fp, fp]
        L.ora_hypothesis_fusion.argtypes = [C.POINTER(OraParams), fp, fp, C.c_int, fp, fp]
        L.ora_stereo_search_constraints.argtypes = [fp, C.c_int, fp, fp]
        L.ora_rot_in_plane_median.argtypes = [vp, fp, C.c_int, vp, fp, C.c_int]
        L.ora_rot_in_plane_median.restype = C.c_float
        L.ora_pair_geometry.argtypes = [fp, fp, fp, fp, fp, fp, fp]
        L.ora_search_range.argtypes = [fp, fp, fp, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, fp, fp]
        L.ora_intra_check_planes.argtypes = [C.POINTER(OraParams), vp, vp, C.c_int, C.c_int]
        L.ora_intra_grow_planes.argtypes = [C.POINTER(OraParams), vp, vp, vp, C.c_int, C.c_int]
        L.ora_pixel_depth.argtypes = [fp, fp, fp, C.c_float, C.c_int, C.c_int]
        L.ora_pixel_depth.restype = C.c_float
        L.ora_epipolar_search.argtypes = [vp, C.c_int, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, C.c_float, fp, fp, fp, fp]
        L.ora_cv_gemm_small_33x31.argtypes = [fp, fp, C.c_double, fp, C.c_double, fp]
        L.ora_cv_gemm_small_33x33.argtypes = [fp, fp, fp]
        L.ora_cv_gemm_small_44x41.argtypes = [fp, fp, fp]
        L.ora_cv_gemm_acc64_ABt.argtypes = [fp, fp, C.c_double, fp]
        L.ora_cv_gemm_acc64_dot.argtypes = [fp, fp, C.c_int, C.c_double]
        L.ora_cv_gemm_acc64_dot.restype = C.c_float
        L.ora_cv_invert33.argtypes = [fp, fp]
        L.ora_cv_solve_lu33.argtypes = [fp, fp, fp]
        _lib = L
    return _lib


def _f(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(C.POINTER(C.c_float))


def _i(a):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a, a.ctypes.data_as(C.POINTER(C.c_int))


def default_params():
    p = OraParams()
    lib().ora_default_params(C.byref(p))
    return p


class Oracle:
    """Key-frame store + the PM.cc stages, CPU."""

    def __init__(self, params=None, threads=1):
        self.L = lib()
        self.h = self.L.ora_create(C.byref(params) if params is not None else None)
        self.shape = {}
        self.set_threads(threads)

    def set_threads(self, n):
        self.L.ora_set_threads(int(n))

    def close(self):
        if self.h:
            self.L.ora_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def add_keyframe(self, kf_id, gray, K, Tcw):
        gray = np.ascontiguousarray(gray, dtype=np.uint8)
        h, w = gray.shape
        _, kp = _f(K)
        _, tp = _f(np.asarray(Tcw).reshape(-1)[:12])
        rc = self.L.ora_add_keyframe(self.h, kf_id, gray.ctypes.data, w, h, w, kp, tp)
        assert rc == 0
        self.shape[kf_id] = (h, w)

    def set_pose(self, kf_id, Tcw):
        _, tp = _f(np.asarray(Tcw).reshape(-1)[:12])
        assert self.L.ora_set_pose(self.h, kf_id, tp) == 0

    def semidense(self, kf_id, nbr_ids, nbr_rot, min_depth, max_depth):
        ids, ip = _i(nbr_ids)
        rot, rp = _f(nbr_rot)
        rc = self.L.ora_semidense(self.h, kf_id, ip, rp, len(ids), float(min_depth), float(max_depth))
        assert rc == 0, rc

    def interkf(self, kf_id, nbr_ids):
        ids, ip = _i(nbr_ids)
        rc = self.L.ora_interkf(self.h, kf_id, ip, len(ids))
        assert rc == 0, rc

    def pointset(self, kf_id):
        assert self.L.ora_pointset(self.h, kf_id) == 0

    def set_maps(self, kf_id, depth, sigma):
        d = np.ascontiguousarray(depth, np.float32)
        s = np.ascontiguousarray(sigma, np.float32)
        assert self.L.ora_set_maps(self.h, kf_id, d.ctypes.data, s.ctypes.data) == 0

    def confident_points(self, kf_id, max_sigma=0.01, max_points=0):
        h, w = self.shape[kf_id]
        cap = h * w
        xyz = np.empty((cap, 3), np.float32)
        sig = np.empty(cap, np.float32)
        pix = np.empty(cap, np.uint32)
        n = self.L.ora_confident_points(self.h, kf_id, float(max_sigma), int(max_points), cap, xyz.ctypes.data, sig.ctypes.data,
                                        pix.ctypes.data)
        assert n >= 0, n
        return xyz[:n].copy(), sig[:n].copy(), pix[:n].copy(), n

    def plane(self, kf_id, what):
        h, w = self.shape[kf_id]
        out = np.empty((h, w * 3 if what == POINTS else w), np.float32)
        rc = self.L.ora_get_plane(self.h, kf_id, what, out.ctypes.data)
        assert rc == 0, rc
        return out

    def istddev(self, kf_id):
        return float(self.L.ora_get_istddev(self.h, kf_id))

    def epipolar_search(self, kf1, kf2, x, y, min_depth, max_depth, rot):
        o = [C.c_float() for _ in range(4)]
        ok = self.L.ora_epipolar_search(self.h, kf1, kf2, x, y, min_depth, max_depth, rot, *[C.byref(v) for v in o])
        return ok, [v.value for v in o]


def intra_planes(depth, sigma, grad=None, grow=False, params=None):
    """IntraKeyFrameDepthChecking (grow=False) / IntraKeyFrameDepthGrowing (grow=True) on host planes."""
    d = np.array(depth, np.float32, order="C")
    s = np.array(sigma, np.float32, order="C")
    h, w = d.shape
    pp = C.byref(params) if params is not None else None
    if grow:
        g = np.ascontiguousarray(grad, np.float32)
        lib().ora_intra_grow_planes(pp, d.ctypes.data, s.ctypes.data, g.ctypes.data, w, h)
    else:
        lib().ora_intra_check_planes(pp, d.ctypes.data, s.ctypes.data, w, h)
    return d, s


def search_range(K1, R21, t21, cols, px, py, mind, maxd):
    a = [_f(x) for x in (K1, np.asarray(R21).reshape(-1), t21)]
    lo, hi = C.c_float(), C.c_float()
    lib().ora_search_range(a[0][1], a[1][1], a[2][1], int(cols), int(px), int(py), float(mind), float(maxd),
                           C.byref(lo), C.byref(hi))
    return lo.value, hi.value


def stereo_search_constraints(depths):
    d, dp = _f(depths)
    mn, mx = C.c_float(), C.c_float()
    lib().ora_stereo_search_constraints(dp, len(d), C.byref(mn), C.byref(mx))
    return mn.value, mx.value


def pair_geometry(K1, Tcw1, K2, Tcw2):
    a = [_f(x) for x in (K1, np.asarray(Tcw1).reshape(-1)[:12], K2, np.asarray(Tcw2).reshape(-1)[:12])]
    Fp, Rp, tp = (C.c_float * 9)(), (C.c_float * 9)(), (C.c_float * 3)()
    lib().ora_pair_geometry(a[0][1], a[1][1], a[2][1], a[3][1], Fp, Rp, tp)
    return np.array(Fp, np.float32).reshape(3, 3), np.array(Rp, np.float32).reshape(3, 3), np.array(tp, np.float32)
