"""The C-ABI shared library: loads on a CPU-only box, exports every symbol include/sdpm.h declares,
its host-side helpers agree with the oracle, and the product fails loudly without a CUDA device."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, "include", "sdpm.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(sd_[a-z_0-9]+)\s*\(", src)))


def test_header_symbols_are_exported(pkg):
    names = header_symbols()
    assert len(names) >= 30
    L = pkg.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in include/sdpm.h but not exported"
    assert sorted(pkg.SYMBOLS) == names, "python binding table and header disagree"
    out = subprocess.run(["nm", "-D", "--defined-only", pkg.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (sd_[a-z_0-9]+)", out))
    assert exported == set(names)


def test_signatures_have_no_foreign_types():
    src = open(os.path.join(ROOT, "include", "sdpm.h")).read()
    for bad in ("torch", "at::", "cv::", "cudaStream_t", "std::"):
        assert bad not in re.sub(r"/\*.*?\*/", "", src, flags=re.S)


def test_library_is_sm100a_only(pkg):
    out = subprocess.run(["cuobjdump", "-lelf", pkg.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_params_defaults_match_reference_defines(pkg):
    p = pkg.default_params()   # PM.h:38-49 and the literals of PM.cc
    assert (p.lambdaG, p.lambdaL, p.lambdaTheta, p.lambdaN) == (8.0, 80.0, 45.0, 3)
    assert (p.theta, p.chi2_fuse, p.chi2_inter, p.slope_limit, p.border) == (0.23, 5.99, 3.84, 4.0, 2)


def test_product_does_not_touch_the_oracle():
    """The product path may not import, link or call anything under oracle/."""
    pk = os.path.join(ROOT, "orb-slam-free-space-carving_b200")
    for dp, _, files in os.walk(pk):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "pm_oracle" not in txt and "oracle/" not in txt and "cvlike" not in txt, os.path.join(dp, f)
    out = subprocess.run(["ldd", os.path.join(pk, "lib", "libsdpm.so")], capture_output=True, text=True).stdout
    assert "oracle" not in out


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="CPU-only check")
def test_context_fails_loudly_without_gpu(pkg):
    with pytest.raises(pkg.SdError) as e:
        pkg.Context(0)
    assert e.value.code == -5   # SD_ERR_NO_DEVICE: there is no CPU fallback


def test_host_stereo_search_constraints_matches_oracle(pkg, po):
    rng = np.random.default_rng(3)
    for n in (1, 2, 17, 1000):
        d = rng.uniform(0.7, 1.4, n).astype(np.float32)
        assert pkg.stereo_search_constraints(d) == po.stereo_search_constraints(d)
    d = np.array([0.1, 2.5, 0.2, 3.0], np.float32)    # mean - 2 sigma < 0: negative min_depth, no guard (PM.cc:382)
    got, ref = pkg.stereo_search_constraints(d), po.stereo_search_constraints(d)
    assert got == ref and got[0] < 0


def test_host_rot_in_plane_matches_oracle(pkg, po):
    rng = np.random.default_rng(4)
    L = po.lib()
    fp = C.POINTER(C.c_float)
    for trial in range(20):
        n1, n2 = rng.integers(0, 300, 2)
        mp1 = rng.integers(-1, 200, n1).astype(np.int64)
        mp2 = rng.integers(-1, 200, n2).astype(np.int64)
        a1 = rng.uniform(-20, 360, n1).astype(np.float32)
        a2 = rng.uniform(-20, 360, n2).astype(np.float32)
        ref = L.ora_rot_in_plane_median(mp1.ctypes.data, a1.ctypes.data_as(fp), int(n1), mp2.ctypes.data, a2.ctypes.data_as(fp), int(n2))
        assert pkg.rot_in_plane_median(mp1, a1, mp2, a2) == ref


def test_host_pair_geometry_matches_oracle(pkg, po, synth):
    """F12 / R21 / t21 (PM.cc:972-987, 890-891): the product's host math and the oracle's are two
    independent restatements of the same OpenCV arithmetic and must agree to the bit."""
    rng = np.random.default_rng(6)
    sc = synth.Scene("720p")
    Ks = [sc.K, np.array([525, 525, 319.5, 239.5], np.float32), np.array([1650, 1650, 959.5, 539.5], np.float32)]
    for trial in range(200):
        if trial < 100:
            k1, k2 = rng.integers(0, 400, 2)
            T1, T2 = sc.pose(int(k1)), sc.pose(int(k2))
        else:
            def rnd():
                q = rng.standard_normal(4); q /= np.linalg.norm(q); w, a, b, c = q
                R = np.array([[1-2*(b*b+c*c), 2*(a*b-c*w), 2*(a*c+b*w)], [2*(a*b+c*w), 1-2*(a*a+c*c), 2*(b*c-a*w)], [2*(a*c-b*w), 2*(b*c+a*w), 1-2*(a*a+b*b)]])
                return np.hstack([R, rng.uniform(-1, 1, (3, 1))]).astype(np.float32).reshape(-1)
            T1, T2 = rnd(), rnd()
        K1, K2 = Ks[trial % 3], Ks[(trial // 3) % 3]
        got, ref = pkg.pair_geometry(K1, T1, K2, T2), po.pair_geometry(K1, T1, K2, T2)
        for g, r in zip(got, ref):
            assert np.array_equal(g, r), trial
