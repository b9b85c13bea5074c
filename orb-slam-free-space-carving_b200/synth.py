"""ctypes binding of the synthetic key-frame generator (synth/scene.cc): input generation for
tests and bench.py.  CPU ray-casting, no graphics API."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "lib", "libsdsynth.so")


class SynthConfig(C.Structure):
    _fields_ = [("w", C.c_int), ("h", C.c_int), ("fx", C.c_float), ("fy", C.c_float), ("cx", C.c_float),
                ("cy", C.c_float), ("seed", C.c_uint32)]


# intrinsics per configuration (720p: chris_logic_HD720.yaml:8-11)
CAMERAS = {
    "480p": dict(w=640, h=480, fx=525.0, fy=525.0, cx=319.5, cy=239.5),
    "720p": dict(w=1280, h=720, fx=1101.107438, fy=1095.367233, cx=604.565580, cy=365.568908),
    "1080p": dict(w=1920, h=1080, fx=1650.0, fy=1650.0, cx=959.5, cy=539.5),
}

_lib = None


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            raise RuntimeError(f"{_LIB_PATH} missing: run `python __graft_entry__.py build`")
        L = C.CDLL(_LIB_PATH)
        fp = C.POINTER(C.c_float)
        L.synth_pose.argtypes = [C.c_int, fp, fp]
        L.synth_render.argtypes = [C.POINTER(SynthConfig), C.c_int, C.c_void_p, C.c_void_p]
        L.synth_sparse_depths.argtypes = [C.POINTER(SynthConfig), C.c_int, C.c_int, fp]
        L.synth_neighbours.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]
        L.synth_neighbours.restype = C.c_int
        L.synth_rot_in_plane.argtypes = [C.c_int, C.c_int]
        L.synth_rot_in_plane.restype = C.c_float
        _lib = L
    return _lib


class Scene:
    """A seeded synthetic sequence: images, ground-truth poses/depth, neighbour lists, ORB stand-ins."""

    def __init__(self, camera="480p", seed=1234, **override):
        cam = dict(CAMERAS[camera]) if isinstance(camera, str) else dict(camera)
        cam.update(override)
        self.cam = cam
        self.cfg = SynthConfig(cam["w"], cam["h"], cam["fx"], cam["fy"], cam["cx"], cam["cy"], seed)
        self.w, self.h = cam["w"], cam["h"]
        self.K = np.array([cam["fx"], cam["fy"], cam["cx"], cam["cy"]], np.float32)

    def pose(self, k):
        T = np.zeros(12, np.float32)
        lib().synth_pose(int(k), T.ctypes.data_as(C.POINTER(C.c_float)), None)
        return T

    def render(self, k, with_depth=False, out=None):
        img = out if out is not None else np.empty((self.h, self.w), np.uint8)
        gt = np.empty((self.h, self.w), np.float32) if with_depth else None
        lib().synth_render(C.byref(self.cfg), int(k), img.ctypes.data, gt.ctypes.data if with_depth else None)
        return (img, gt) if with_depth else img

    def sparse_depths(self, k, n=1000):
        d = np.empty(n, np.float32)
        lib().synth_sparse_depths(C.byref(self.cfg), int(k), n, d.ctypes.data_as(C.POINTER(C.c_float)))
        return d

    @staticmethod
    def neighbours(k, nkf, n):
        ids = (C.c_int * n)()
        got = lib().synth_neighbours(int(k), int(nkf), int(n), ids)
        return np.array(ids[:got], np.int32)

    @staticmethod
    def rot_in_plane(k1, k2):
        return float(lib().synth_rot_in_plane(int(k1), int(k2)))
