"""B200-native semi-dense probability mapping (the ProbabilityMapping hot path of
atlas-jj/ORB-SLAM-free-space-carving).

The product is `lib/libsdpm.so`: hand-written CUDA kernels for sm_100a behind the C ABI declared in
`include/sdpm.h`; the C++ host mirror of the reference class lives in `host/`.  This Python
module is only a ctypes binding of that C ABI for the test-suite and bench.py.  There is no CPU
fallback: importing works anywhere, but `Context()` raises unless the library is built and a
CUDA device is present.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libsdpm.so")
INCLUDE_DIR = os.path.join(os.path.dirname(_HERE), "include")

SD_MAX_NEIGHBOURS = 40
SD_STATIC_MAX_PIXELS = 1280 * 720   # include/sdpm.h: largest plane sd_intra_depth_check_planes_static accepts
GRAD, THETA, DEPTH, SIGMA, DEPTH_CHECKED, POINTS = range(6)

STATUS = {0: "SD_OK", -1: "SD_ERR_INVALID_ARG", -2: "SD_ERR_UNKNOWN_KEYFRAME", -3: "SD_ERR_NOT_READY",
          -4: "SD_ERR_CUDA", -5: "SD_ERR_NO_DEVICE", -6: "SD_ERR_SHAPE"}


class SdError(RuntimeError):
    def __init__(self, code, msg=""):
        super().__init__(f"{STATUS.get(code, code)}: {msg}")
        self.code = code


class SdParams(C.Structure):
    _fields_ = [("lambdaG", C.c_float), ("lambdaL", C.c_float), ("lambdaTheta", C.c_float), ("lambdaN", C.c_int),
                ("theta", C.c_double), ("chi2_fuse", C.c_double), ("chi2_inter", C.c_double),
                ("slope_limit", C.c_float), ("border", C.c_int), ("interkf_inplace", C.c_int)]


class SdLoopItem(C.Structure):
    _fields_ = [("kf_id", C.c_int), ("nbr_ids", C.POINTER(C.c_int)), ("nbr_rot_deg", C.POINTER(C.c_float)), ("n", C.c_int),
                ("min_invdepth", C.c_float), ("max_invdepth", C.c_float), ("finish_only", C.c_int), ("download", C.c_int),
                ("out_depth", C.c_void_p), ("out_sigma", C.c_void_p), ("out_points", C.c_void_p),
                ("out_records", C.c_void_p), ("capacity", C.c_int), ("out_count", C.POINTER(C.c_int))]


class SdHaloPeer(C.Structure):
    _fields_ = [("peer", C.c_int), ("n_send", C.c_int), ("send_ids", C.POINTER(C.c_int)), ("n_recv", C.c_int),
                ("recv_ids", C.POINTER(C.c_int))]


SD_LOOP_INTERKF = 1
SD_DL_NONE, SD_DL_DENSE, SD_DL_DENSE_POINTS, SD_DL_COMPACT, SD_DL_COMPACT_POINTS = range(5)
# records of sd_download_compact (include/sdpm.h: SdCompactDepth / SdCompactPoint)
COMPACT_DEPTH = np.dtype([("pixel", "<u4"), ("depth", "<f4"), ("sigma", "<f4")])
COMPACT_POINT = np.dtype([("pixel", "<u4"), ("depth", "<f4"), ("sigma", "<f4"), ("x", "<f4"), ("y", "<f4"), ("z", "<f4")])


class SdDevicePlanes(C.Structure):
    _fields_ = [("width", C.c_int), ("height", C.c_int), ("pitch_f32", C.c_int), ("pitch_u8", C.c_int),
                ("gray", C.c_void_p), ("grad", C.c_void_p), ("theta", C.c_void_p), ("depth", C.c_void_p),
                ("sigma", C.c_void_p), ("depth_checked", C.c_void_p), ("points", C.c_void_p)]


_lib = None
_fp, _ip, _vp = C.POINTER(C.c_float), C.POINTER(C.c_int), C.c_void_p

# every symbol include/sdpm.h declares: (restype, argtypes)
SYMBOLS = {
    "sd_default_params": (C.c_int, [C.POINTER(SdParams)]),
    "sd_create": (C.c_int, [C.c_int, C.POINTER(SdParams), C.POINTER(_vp)]),
    "sd_destroy": (C.c_int, [_vp]),
    "sd_upload_keyframe": (C.c_int, [_vp, C.c_int, _vp, C.c_int, C.c_int, C.c_int, _fp, _fp]),
    "sd_upload_keyframe_device": (C.c_int, [_vp, C.c_int, _vp, C.c_int, C.c_int, C.c_int, _fp, _fp]),
    "sd_set_pose": (C.c_int, [_vp, C.c_int, _fp]),
    "sd_release_keyframe": (C.c_int, [_vp, C.c_int]),
    "sd_semidense": (C.c_int, [_vp, C.c_int, _ip, _fp, C.c_int, C.c_float, C.c_float]),
    "sd_interkf_check": (C.c_int, [_vp, C.c_int, _ip, C.c_int]),
    "sd_update_pointset": (C.c_int, [_vp, C.c_int]),
    "sd_interkf_check_points": (C.c_int, [_vp, C.c_int, _ip, C.c_int]),
    "sd_semidense_loop": (C.c_int, [_vp, C.POINTER(SdLoopItem), C.c_int, C.c_int, _ip]),
    "sd_download_compact": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, _vp, _ip]),
    "sd_download_compact_async": (C.c_int, [_vp, C.c_int, C.c_int, C.c_int, _vp, _ip]),
    "sd_confident_points_fetch": (C.c_int, [_vp, C.c_int, _vp, _vp, _vp, _ip]),
    "sd_import_maps_host": (C.c_int, [_vp, C.c_int, _vp, _vp]),
    "sd_nccl_unique_id": (C.c_int, [_vp]),
    "sd_comm_init": (C.c_int, [_vp, C.c_int, C.c_int, _vp]),
    "sd_comm_destroy": (C.c_int, [_vp]),
    "sd_halo_exchange": (C.c_int, [_vp, C.POINTER(SdHaloPeer), C.c_int]),
    "sd_confident_points": (C.c_int, [_vp, C.c_int, C.c_double, C.c_int, C.c_int, _vp, _vp, _vp, _ip]),
    "sd_confident_points_device": (C.c_int, [_vp, C.c_int, C.c_double, C.c_int, C.POINTER(_vp), C.POINTER(_vp), C.POINTER(_vp), _ip, _ip]),
    "sd_download": (C.c_int, [_vp, C.c_int, _vp, _vp, _vp]),
    "sd_download_async": (C.c_int, [_vp, C.c_int, _vp, _vp, _vp]),
    "sd_download_plane": (C.c_int, [_vp, C.c_int, C.c_int, _vp]),
    "sd_get_istddev": (C.c_int, [_vp, C.c_int, _fp]),
    "sd_get_active_count": (C.c_int, [_vp, C.c_int, _ip]),
    "sd_device_ptrs": (C.c_int, [_vp, C.c_int, C.POINTER(SdDevicePlanes)]),
    "sd_import_maps": (C.c_int, [_vp, C.c_int, _vp, _vp]),
    "sd_export_maps": (C.c_int, [_vp, C.c_int, _vp, _vp]),
    "sd_alloc_host": (C.c_int, [_vp, C.c_uint64, C.POINTER(_vp)]),
    "sd_free_host": (C.c_int, [_vp, _vp]),
    "sd_download_fence": (C.c_int, [_vp, C.POINTER(C.c_uint64)]),
    "sd_download_wait": (C.c_int, [_vp, C.c_uint64]),
    "sd_sync": (C.c_int, [_vp]),
    "sd_last_error": (C.c_char_p, [_vp]),
    "sd_compute_stream": (C.c_int, [_vp, C.POINTER(_vp)]),
    "sd_launch_count": (C.c_int, [_vp, C.POINTER(C.c_uint64)]),
    "sd_span_begin": (C.c_int, [_vp]),
    "sd_span_end": (C.c_int, [_vp, C.POINTER(C.c_double)]),
    "sd_timing_enable": (C.c_int, [_vp, C.c_int]),
    "sd_timing_get": (C.c_int, [_vp, C.c_char_p, C.POINTER(C.c_double), C.POINTER(C.c_uint64)]),
    "sd_timing_reset": (C.c_int, [_vp]),
    "sd_stereo_search_constraints": (C.c_int, [_fp, C.c_int, _fp, _fp]),
    "sd_rot_in_plane_median": (C.c_float, [_vp, _fp, C.c_int, _vp, _fp, C.c_int]),
    "sd_pair_geometry": (C.c_int, [_fp, _fp, _fp, _fp, _fp, _fp, _fp]),
    "sd_search_range": (C.c_int, [_fp, _fp, _fp, C.c_int, C.c_int, C.c_int, C.c_float, C.c_float, _fp, _fp]),
    "sd_intra_depth_check_planes": (C.c_int, [_vp, _vp, _vp, _vp, C.c_int, C.c_int]),
    "sd_intra_depth_grow_planes": (C.c_int, [_vp, _vp, _vp, _vp, C.c_int, C.c_int]),
    "sd_intra_depth_check_planes_static": (C.c_int, [_vp, _vp, _vp, _vp, C.c_int, C.c_int]),
    "sd_probe_epipolar": (C.c_int, [_vp, C.c_int, C.c_int, _ip, C.c_int, C.c_float, C.c_float, C.c_float, _fp]),
    "sd_probe_fusion": (C.c_int, [_vp, _fp, _fp, C.c_int, _fp, _fp, _ip]),
}


def lib():
    """Load lib/libsdpm.so.  Fails loudly when it has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing -- build it with `python __graft_entry__.py build` "
                               "(there is no CPU fallback for the CUDA path)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SYMBOLS.items():
            fn = getattr(L, name)  # AttributeError if the library does not export a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _f(a):
    a = np.ascontiguousarray(a, dtype=np.float32)
    return a, a.ctypes.data_as(_fp)


def _i(a):
    a = np.ascontiguousarray(a, dtype=np.int32)
    return a, a.ctypes.data_as(_ip)


def default_params():
    p = SdParams()
    lib().sd_default_params(C.byref(p))
    return p


def nccl_unique_id():
    """128 bytes from ncclGetUniqueId (rank 0 calls this and distributes them)."""
    buf = C.create_string_buffer(128)
    rc = lib().sd_nccl_unique_id(buf)
    if rc:
        raise SdError(rc, "sd_nccl_unique_id (libnccl.so.2 not loadable?)")
    return buf.raw


def expand_compact(records, h, w, points=False):
    """Dense (depth, sigma[, points]) planes from compact records: what ProbabilityMapping::FetchResults does in C++."""
    d = np.zeros((h, w), np.float32)
    s = np.zeros((h, w), np.float32)
    y, x = (records["pixel"] >> 16).astype(np.int64), (records["pixel"] & 0xffff).astype(np.int64)
    d[y, x] = records["depth"]
    s[y, x] = records["sigma"]
    if not points:
        return d, s
    p = np.zeros((h, 3 * w), np.float32)
    p[y, 3 * x], p[y, 3 * x + 1], p[y, 3 * x + 2] = records["x"], records["y"], records["z"]
    return d, s, p


def stereo_search_constraints(depths):
    """StereoSearchConstraints (PM.cc:370-383): returns (min_depth, max_depth) = inverse-depth interval."""
    d, dp = _f(depths)
    mn, mx = C.c_float(), C.c_float()
    rc = lib().sd_stereo_search_constraints(dp, len(d), C.byref(mn), C.byref(mx))
    if rc:
        raise SdError(rc)
    return mn.value, mx.value


def rot_in_plane_median(mp1, ang1, mp2, ang2):
    m1 = np.ascontiguousarray(mp1, np.int64)
    m2 = np.ascontiguousarray(mp2, np.int64)
    a1, a1p = _f(ang1)
    a2, a2p = _f(ang2)
    return float(lib().sd_rot_in_plane_median(m1.ctypes.data, a1p, len(m1), m2.ctypes.data, a2p, len(m2)))


def pair_geometry(K1, Tcw1, K2, Tcw2):
    a = [_f(x) for x in (K1, np.asarray(Tcw1).reshape(-1)[:12], K2, np.asarray(Tcw2).reshape(-1)[:12])]
    F, R, t = (C.c_float * 9)(), (C.c_float * 9)(), (C.c_float * 3)()
    rc = lib().sd_pair_geometry(a[0][1], a[1][1], a[2][1], a[3][1], F, R, t)
    if rc:
        raise SdError(rc)
    return np.array(F, np.float32).reshape(3, 3), np.array(R, np.float32).reshape(3, 3), np.array(t, np.float32)


def search_range(K1, Tcw1, Tcw2, cols, px, py, mind, maxd):
    """GetSearchRange (PM.cc:877-910): (umin, umax) clamped to [0, cols]."""
    a = [_f(x) for x in (K1, np.asarray(Tcw1).reshape(-1)[:12], np.asarray(Tcw2).reshape(-1)[:12])]
    lo, hi = C.c_float(), C.c_float()
    rc = lib().sd_search_range(a[0][1], a[1][1], a[2][1], int(cols), int(px), int(py), float(mind), float(maxd),
                               C.byref(lo), C.byref(hi))
    if rc:
        raise SdError(rc)
    return lo.value, hi.value


class Context:
    """One sd_ctx (one GPU).  Thin wrapper: every method is one C-ABI call."""

    def __init__(self, device=0, params=None):
        self.L = lib()
        h = _vp()
        rc = self.L.sd_create(int(device), C.byref(params) if params is not None else None, C.byref(h))
        if rc:
            raise SdError(rc, "sd_create failed (no usable CUDA device?)")
        self.h = h
        self.shape = None

    def _ck(self, rc):
        if rc:
            raise SdError(rc, (self.L.sd_last_error(self.h) or b"").decode())

    def close(self):
        if getattr(self, "h", None):
            self.L.sd_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def upload_keyframe(self, kf_id, gray, K, Tcw):
        """gray: HxW uint8 numpy array (or any object with .ctypes / int pointer via upload_keyframe_ptr)."""
        gray = np.ascontiguousarray(gray, dtype=np.uint8)
        h, w = gray.shape
        self.upload_keyframe_ptr(kf_id, gray.ctypes.data, w, h, w, K, Tcw)

    def upload_keyframe_ptr(self, kf_id, ptr, w, h, stride, K, Tcw, device=False):
        _, kp = _f(K)
        _, tp = _f(np.asarray(Tcw).reshape(-1)[:12])
        fn = self.L.sd_upload_keyframe_device if device else self.L.sd_upload_keyframe
        self._ck(fn(self.h, int(kf_id), ptr, w, h, stride, kp, tp))
        self.shape = (h, w)

    def set_pose(self, kf_id, Tcw):
        _, tp = _f(np.asarray(Tcw).reshape(-1)[:12])
        self._ck(self.L.sd_set_pose(self.h, int(kf_id), tp))

    def release_keyframe(self, kf_id):
        self._ck(self.L.sd_release_keyframe(self.h, int(kf_id)))

    def semidense(self, kf_id, nbr_ids, nbr_rot, min_depth, max_depth):
        ids, ip = _i(nbr_ids)
        rot, rp = _f(nbr_rot)
        self._ck(self.L.sd_semidense(self.h, int(kf_id), ip, rp, len(ids), float(min_depth), float(max_depth)))

    def interkf_check(self, kf_id, nbr_ids):
        ids, ip = _i(nbr_ids)
        self._ck(self.L.sd_interkf_check(self.h, int(kf_id), ip, len(ids)))

    def update_pointset(self, kf_id):
        self._ck(self.L.sd_update_pointset(self.h, int(kf_id)))

    def interkf_check_points(self, kf_id, nbr_ids):
        """InterKeyFrameDepthChecking + UpdateSemiDensePointSet (PM.cc:302-304) in one launch."""
        ids, ip = _i(nbr_ids)
        self._ck(self.L.sd_interkf_check_points(self.h, int(kf_id), ip, len(ids)))

    def make_loop(self, items, interkf=True):
        """Marshal a key-frame list for sd_semidense_loop once (the arrays stay alive with the returned object).
        items: dicts with kf_id, nbrs, rots, prior=(min, max) and optional download=(mode, ...) entries:
        (SD_DL_DENSE, depth_ptr, sigma_ptr), (SD_DL_DENSE_POINTS, depth_ptr, sigma_ptr, points_ptr),
        (SD_DL_COMPACT | SD_DL_COMPACT_POINTS, records_ptr, capacity, count_ptr)."""
        arr = (SdLoopItem * len(items))()
        keep = []
        for a, it in zip(arr, items):
            ids, ip = _i(it["nbrs"])
            rot, rp = _f(it["rots"])
            keep += [ids, rot]
            a.kf_id, a.nbr_ids, a.nbr_rot_deg, a.n = int(it["kf_id"]), ip, rp, len(ids)
            a.min_invdepth, a.max_invdepth = float(it["prior"][0]), float(it["prior"][1])
            a.finish_only = 1 if it.get("finish_only") else 0
            dl = it.get("download")
            a.download = SD_DL_NONE
            if dl:
                a.download = int(dl[0])
                if dl[0] in (SD_DL_DENSE, SD_DL_DENSE_POINTS):
                    a.out_depth, a.out_sigma = dl[1], dl[2]
                    a.out_points = dl[3] if dl[0] == SD_DL_DENSE_POINTS else None
                else:
                    a.out_records, a.capacity = dl[1], int(dl[2])
                    a.out_count = C.cast(dl[3], C.POINTER(C.c_int)) if dl[3] else None
        done = np.zeros(len(items), np.int32)
        return {"arr": arr, "n": len(items), "flags": SD_LOOP_INTERKF if interkf else 0, "keep": keep, "done": done}

    def semidense_loop(self, loop):
        """SemiDenseLoop over a key-frame list in one C call (PM.cc:137-307); returns the per-item inter-KF flags."""
        self._ck(self.L.sd_semidense_loop(self.h, loop["arr"], loop["n"], loop["flags"], loop["done"].ctypes.data_as(_ip)))
        return loop["done"]

    def download_compact(self, kf_id, points=False, capacity=None):
        """Records (numpy structured array, COMPACT_DEPTH / COMPACT_POINT) of the pixels that carry a depth or a sigma."""
        h, w = self.shape
        cap = int(capacity) if capacity is not None else (w - 4) * (h - 4)
        rec = np.empty(cap, COMPACT_POINT if points else COMPACT_DEPTH)
        n = C.c_int()
        self._ck(self.L.sd_download_compact(self.h, int(kf_id), 1 if points else 0, cap, rec.ctypes.data, C.byref(n)))
        return rec[:min(n.value, cap)], n.value

    def download_compact_async_ptr(self, kf_id, points, capacity, records_ptr, count_ptr):
        self._ck(self.L.sd_download_compact_async(self.h, int(kf_id), 1 if points else 0, int(capacity), records_ptr,
                                                  C.cast(count_ptr, _ip) if count_ptr else None))

    def import_maps_host(self, kf_id, depth, sigma):
        d = np.ascontiguousarray(depth, np.float32)
        s = np.ascontiguousarray(sigma, np.float32)
        self._ck(self.L.sd_import_maps_host(self.h, int(kf_id), d.ctypes.data, s.ctypes.data))

    # ---- multi-GPU: NCCL halo exchange behind the C ABI ----
    def comm_init(self, world, rank, unique_id: bytes):
        buf = C.create_string_buffer(bytes(unique_id), 128)
        self._ck(self.L.sd_comm_init(self.h, int(world), int(rank), buf))

    def comm_destroy(self):
        self._ck(self.L.sd_comm_destroy(self.h))

    def make_halo(self, send, recv):
        """send / recv: {peer: [key-frame ids]} -> marshalled SdHaloPeer array (kept alive by the returned object)."""
        peers = sorted(set(send) | set(recv))
        arr = (SdHaloPeer * max(len(peers), 1))()
        keep = []
        for a, p in zip(arr, peers):
            si, sp = _i(sorted(send.get(p, [])))
            ri, rp = _i(sorted(recv.get(p, [])))
            keep += [si, ri]
            a.peer, a.n_send, a.send_ids, a.n_recv, a.recv_ids = int(p), len(si), sp, len(ri), rp
        return {"arr": arr, "n": len(peers), "keep": keep}

    def halo_exchange(self, halo):
        self._ck(self.L.sd_halo_exchange(self.h, halo["arr"], halo["n"]))

    def confident_points(self, kf_id, max_sigma=0.01, max_points=0, capacity=None):
        """(xyz [n,3], sigma [n], pixel [n] = y << 16 | x, count) of the pixels PM.cc:120-121 keeps; see sd_confident_points."""
        h, w = self.shape
        cap = int(capacity) if capacity is not None else (int(max_points) if 0 < max_points < h * w else h * w)
        xyz = np.empty((cap, 3), np.float32)
        sig = np.empty(cap, np.float32)
        pix = np.empty(cap, np.uint32)
        n = C.c_int()
        self._ck(self.L.sd_confident_points(self.h, int(kf_id), float(max_sigma), int(max_points), cap, xyz.ctypes.data,
                                            sig.ctypes.data, pix.ctypes.data, C.byref(n)))
        m = min(n.value, cap)
        return xyz[:m], sig[:m], pix[:m], n.value

    def confident_points_device(self, kf_id, max_sigma=0.01, max_points=0):
        """Device pointers (xyz, sigma, pixel) into the context's scratch + (selected, confident) counts."""
        px, ps, pp = _vp(), _vp(), _vp()
        n, nc = C.c_int(), C.c_int()
        self._ck(self.L.sd_confident_points_device(self.h, int(kf_id), float(max_sigma), int(max_points), C.byref(px), C.byref(ps),
                                                   C.byref(pp), C.byref(n), C.byref(nc)))
        return px.value, ps.value, pp.value, n.value, nc.value

    def download(self, kf_id, points=False):
        h, w = self.shape
        d = np.empty((h, w), np.float32)
        s = np.empty((h, w), np.float32)
        p = np.empty((h, 3 * w), np.float32) if points else None
        self._ck(self.L.sd_download(self.h, int(kf_id), d.ctypes.data, s.ctypes.data, p.ctypes.data if points else None))
        return (d, s, p) if points else (d, s)

    def download_async_ptr(self, kf_id, depth_ptr, sigma_ptr, points_ptr=None):
        self._ck(self.L.sd_download_async(self.h, int(kf_id), depth_ptr, sigma_ptr, points_ptr))

    def plane(self, kf_id, which):
        h, w = self.shape
        out = np.empty((h, 3 * w if which == POINTS else w), np.float32)
        self._ck(self.L.sd_download_plane(self.h, int(kf_id), which, out.ctypes.data))
        return out

    def istddev(self, kf_id):
        v = C.c_float()
        self._ck(self.L.sd_get_istddev(self.h, int(kf_id), C.byref(v)))
        return v.value

    def active_count(self, kf_id):
        v = C.c_int()
        self._ck(self.L.sd_get_active_count(self.h, int(kf_id), C.byref(v)))
        return v.value

    def device_ptrs(self, kf_id):
        p = SdDevicePlanes()
        self._ck(self.L.sd_device_ptrs(self.h, int(kf_id), C.byref(p)))
        return p

    def import_maps(self, kf_id, depth_dev_ptr, sigma_dev_ptr):
        self._ck(self.L.sd_import_maps(self.h, int(kf_id), depth_dev_ptr, sigma_dev_ptr))

    def export_maps(self, kf_id, depth_dev_ptr, sigma_dev_ptr):
        self._ck(self.L.sd_export_maps(self.h, int(kf_id), depth_dev_ptr, sigma_dev_ptr))

    def sync(self):
        self._ck(self.L.sd_sync(self.h))

    def compute_stream(self):
        s = _vp()
        self._ck(self.L.sd_compute_stream(self.h, C.byref(s)))
        return s.value

    def launch_count(self):
        v = C.c_uint64()
        self._ck(self.L.sd_launch_count(self.h, C.byref(v)))
        return v.value

    def span_begin(self):
        self._ck(self.L.sd_span_begin(self.h))

    def span_end(self):
        ms = C.c_double()
        self._ck(self.L.sd_span_end(self.h, C.byref(ms)))
        return ms.value

    def timing_enable(self, on=True, serial=False):
        self._ck(self.L.sd_timing_enable(self.h, (2 if serial else 1) if on else 0))

    def timing_reset(self):
        self._ck(self.L.sd_timing_reset(self.h))

    def timing(self):
        out = {}
        for name in ("grad", "search", "fuse", "intra", "inter", "points"):
            ms, n = C.c_double(), C.c_uint64()
            self._ck(self.L.sd_timing_get(self.h, name.encode(), C.byref(ms), C.byref(n)))
            out[name] = (ms.value, n.value)
        return out

    def intra_planes(self, depth, sigma, grad, grow=False):
        """IntraKeyFrameDepthChecking / ...Growing on host planes; returns new (depth, sigma)."""
        d = np.array(depth, np.float32, order="C")
        s = np.array(sigma, np.float32, order="C")
        g = np.ascontiguousarray(grad, np.float32)
        h, w = d.shape
        fn = self.L.sd_intra_depth_grow_planes if grow else self.L.sd_intra_depth_check_planes
        self._ck(fn(self.h, d.ctypes.data, s.ctypes.data, g.ctypes.data, w, h))
        return d, s

    def intra_check_planes_static(self, depth, sigma, grad):
        """sd_intra_depth_check_planes_static: IntraKeyFrameDepthChecking on host planes without any device allocation
        (working planes built into the library, at most STATIC_MAX_PIXELS pixels); returns new (depth, sigma)."""
        d = np.array(depth, np.float32, order="C")
        s = np.array(sigma, np.float32, order="C")
        g = np.ascontiguousarray(grad, np.float32)
        h, w = d.shape
        self._ck(self.L.sd_intra_depth_check_planes_static(self.h, d.ctypes.data, s.ctypes.data, g.ctypes.data, w, h))
        return d, s

    def probe_epipolar(self, kf1, kf2, xy, min_depth, max_depth, rot):
        xy, xp = _i(np.asarray(xy).reshape(-1))
        n = len(xy) // 2
        out = np.empty((n, 6), np.float32)
        self._ck(self.L.sd_probe_epipolar(self.h, int(kf1), int(kf2), xp, n, float(min_depth), float(max_depth), float(rot),
                                          out.ctypes.data_as(_fp)))
        return out

    def probe_fusion(self, depth, sigma):
        d, dp = _f(depth)
        s, sp = _f(sigma)
        od, os_, sup = C.c_float(), C.c_float(), C.c_int()
        self._ck(self.L.sd_probe_fusion(self.h, dp, sp, len(d), C.byref(od), C.byref(os_), C.byref(sup)))
        return od.value, os_.value, bool(sup.value)
