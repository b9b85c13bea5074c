"""The C++ host mirror of the reference class (orb-slam-free-space-carving_b200/host/ProbabilityMapping.{h,cc})
driven through its own test binary, plus the C-ABI entries it forwards to that no other test covers
(sd_search_range on the host; the intra-key-frame stencils on caller-owned planes).

CPU (`not gpu`): the class fails loudly without a device; sd_search_range against the oracle.
GPU: ProbabilityMapping::Run() over a synthetic map == the oracle, bit for bit; the single-pixel
methods; IntraKeyFrameDepthChecking / ...Growing on crafted planes (including planes on which the
growing stage -- a no-op inside the pipeline, SURVEY H7 -- does grow)."""
import os
import struct
import subprocess

import numpy as np
import pytest

from helpers import SMALL_CAM

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
DRIVER = os.path.join(ROOT, "orb-slam-free-space-carving_b200", "lib", "test_pm_class")


def test_class_fails_loudly_without_device(built):
    r = subprocess.run([DRIVER, "--no-device-check"], capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    assert "construction failed loudly" in r.stdout or "a CUDA device is present" in r.stdout


def test_search_range_matches_oracle(pkg, po, synth):
    """GetSearchRange, PM.cc:877-910 (host code in the reference as well)."""
    scene = synth.Scene("480p", seed=5)
    T1, T2 = scene.pose(3), scene.pose(6)
    _, R21, t21 = po.pair_geometry(scene.K, T1, scene.K, T2)
    rng = np.random.default_rng(0)
    for _ in range(300):
        px, py = int(rng.integers(0, scene.w)), int(rng.integers(0, scene.h))
        mind, maxd = float(np.float32(rng.uniform(0.9, 3.0))), float(np.float32(rng.uniform(0.2, 0.9)))
        got = pkg.search_range(scene.K, T1, T2, scene.w, px, py, mind, maxd)
        want = po.search_range(scene.K, R21, t21, scene.w, px, py, mind, maxd)
        assert got == want, (px, py, mind, maxd, got, want)
    # clamping: a huge inverse-depth interval saturates at [0, cols]
    lo, hi = pkg.search_range(scene.K, T1, T2, scene.w, 10, 10, 50.0, 1e-3)
    assert 0.0 <= lo <= hi <= float(scene.w)


def _write_scene(path, scene, plan, images, covis_n, probes, rng):
    nkf = len(plan.ids)
    index = {k: i for i, k in enumerate(plan.ids)}
    with open(path, "wb") as f:
        f.write(struct.pack("<4i", nkf, scene.w, scene.h, covis_n))
        for k in plan.ids:
            f.write(struct.pack("<i", k))
            f.write(np.asarray(scene.K, np.float32).tobytes())
            f.write(np.asarray(plan.poses[k], np.float32).tobytes())
            cov = [index[int(j)] for j in plan.nbrs[k]]
            f.write(struct.pack("<i", len(cov)))
            f.write(np.asarray(cov, np.int32).tobytes())
            d = scene.sparse_depths(k, 1000)
            f.write(struct.pack("<i", len(d)))
            f.write(d.tobytes())
            # key points: every key frame observes the same 50 map points with the in-plane angle of
            # the synthetic trajectory, so that GetRotInPlane's median == Scene.rot_in_plane
            mp = np.arange(50, dtype=np.int64)
            ang = np.full(50, 100.0, np.float32) + np.float32(synth_roll(k))
            f.write(struct.pack("<i", len(mp)))
            f.write(mp.tobytes())
            f.write(ang.tobytes())
            f.write(np.ascontiguousarray(images[k]).tobytes())
        f.write(struct.pack("<i", len(probes)))
        for (i1, i2, x, y, mind, maxd, rot) in probes:
            f.write(struct.pack("<4i3f", i1, i2, x, y, mind, maxd, rot))


_ROLL = {}


def synth_roll(k):
    return _ROLL[k]


def _host_scene(tmp_path, synth, pipeline, nkf=12, n=5, seed=21, with_probes=True):
    scene = synth.Scene(SMALL_CAM, seed=seed)
    plan = pipeline.SequencePlan(scene, nkf, n)
    images = {k: scene.render(k) for k in plan.ids}
    # angles such that angle2 - angle1 == rot_in_plane(k1, k2) is not guaranteed in float; use what the
    # class will compute: rot = (100 + roll_j) - (100 + roll_k), evaluated in float32 like the class does
    for k in plan.ids:
        _ROLL[k] = float(np.float32(synth.Scene.rot_in_plane(0, k)))
    rots = {k: np.array([np.float32(np.float32(100.0) + np.float32(_ROLL[int(j)])) -
                         np.float32(np.float32(100.0) + np.float32(_ROLL[k])) for j in plan.nbrs[k]], np.float32)
            for k in plan.ids}
    probes = []
    if with_probes:
        ys, xs = np.nonzero(np.abs(np.diff(images[plan.ids[4]].astype(np.int32), axis=1)) > 12)
        mind, maxd = plan.prior[plan.ids[4]]
        probes = [(4, 5, int(xs[i]) + 2, int(ys[i]), mind, maxd, 0.0) for i in range(0, len(xs), max(1, len(xs) // 40))][:40]
        probes = [p for p in probes if 2 <= p[2] < scene.w - 2 and 2 <= p[3] < scene.h - 2]
    scene_path = tmp_path / "scene.bin"
    _write_scene(scene_path, scene, plan, images, n, probes, None)
    return scene, plan, images, rots, probes, scene_path


def _oracle_run(po, scene, plan, images, rots):
    ora = po.Oracle(threads=po.lib().ora_max_threads())
    for k in plan.ids:
        ora.add_keyframe(k, images[k], scene.K, plan.poses[k])
    for k in plan.ids:
        ora.semidense(k, plan.nbrs[k], rots[k], *plan.prior[k])
    for k in plan.ids:
        ora.interkf(k, plan.nbrs[k])
        ora.pointset(k)
    return ora


def _check_maps(raw, scene, plan, ora, po):
    """The per-key-frame part of out.bin against the oracle, bit for bit; returns (offset, planes)."""
    px = scene.w * scene.h
    off = 0
    planes = {}
    for k in plan.ids:
        flags = raw[off:off + 8].view(np.int32); off += 8
        assert list(flags) == [1, 1], f"kf {k}: flags {list(flags)}"
        depth = raw[off:off + 4 * px].view(np.float32).reshape(scene.h, scene.w); off += 4 * px
        sigma = raw[off:off + 4 * px].view(np.float32).reshape(scene.h, scene.w); off += 4 * px
        hp = int(raw[off:off + 4].view(np.int32)[0]); off += 4
        assert hp == 1
        pts = raw[off:off + 12 * px].view(np.float32).reshape(scene.h, 3 * scene.w); off += 12 * px
        assert np.array_equal(depth, ora.plane(k, po.DEPTH_CHECKED)), f"depth kf {k}"
        assert np.array_equal(sigma, ora.plane(k, po.SIGMA)), f"sigma kf {k}"
        assert np.array_equal(pts, ora.plane(k, po.POINTS)), f"points kf {k}"
        planes[k] = (depth, sigma)
    return off, planes


@pytest.mark.gpu
def test_class_run_matches_oracle(tmp_path, pkg, po, synth, pipeline):
    scene, plan, images, rots, probes, scene_path = _host_scene(tmp_path, synth, pipeline)
    out_path, obj_path = tmp_path / "out.bin", tmp_path / "cloud.obj"
    r = subprocess.run([DRIVER, str(scene_path), str(out_path), str(obj_path), "0.05"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    ora = _oracle_run(po, scene, plan, images, rots)
    raw = np.fromfile(out_path, np.uint8)
    off, planes = _check_maps(raw, scene, plan, ora, po)
    n_obj = sum(int(((s.astype(np.float64) <= 0.05) & (d > 0.000001)).sum()) for d, s in planes.values())
    assert n_obj > 0
    with open(obj_path) as f:
        lines = f.read().splitlines()
    assert len(lines) == n_obj and all(line.startswith("v ") for line in lines[:50])   # PM.cc:100-132
    expected = []
    for k in plan.ids:   # the device-side selection (sd_confident_points) writes what the oracle's PM.cc:114-127 loop selects
        xyz = ora.confident_points(k, 0.05, 0)[0]
        expected += ["v %f %f %f" % (float(p[0]), float(p[1]), float(p[2])) for p in xyz]
    assert lines == expected
    nfound = 0
    for (i1, i2, x, y, mind, maxd, rot) in probes:
        rec = raw[off:off + 28].view(np.float32); off += 28
        ok, (d, s, bu, bv) = ora.epipolar_search(plan.ids[i1], plan.ids[i2], x, y, mind, maxd, rot)
        assert bool(rec[4]) == bool(ok)
        if ok:
            nfound += 1
            assert rec[0] == np.float32(d) and rec[1] == np.float32(s)
            assert rec[2] == np.float32(bu) and rec[3] == np.float32(bv)        # the refined u*, v* (PM.cc:459-460)
        _, R21, t21 = po.pair_geometry(scene.K, plan.poses[plan.ids[i1]], scene.K, plan.poses[plan.ids[i2]])
        assert (rec[5], rec[6]) == po.search_range(scene.K, R21, t21, scene.w, x, y, mind, maxd)
    assert off == raw.size
    assert nfound > 0


@pytest.mark.gpu
@pytest.mark.parametrize("limit", [0, 14])
def test_class_streamed_equals_batch(tmp_path, pkg, po, synth, pipeline, limit):
    """The replay stub of the live system: the worker thread runs ProbabilityMapping::Run()
    (PM.cc:65-134) while key frames are inserted one at a time; the streamed result is the batch result, also when
    only `limit` key frames may stay resident on the device (finished ones are released as the map grows)."""
    scene, plan, images, rots, _, scene_path = _host_scene(tmp_path, synth, pipeline, nkf=24, n=5, seed=33, with_probes=False)
    out_path, obj_path = tmp_path / "out.bin", tmp_path / "cloud.obj"
    flag = "--stream" + (f"={limit}" if limit else "")
    r = subprocess.run([DRIVER, flag, str(scene_path), str(out_path), str(obj_path)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "stream:" in r.stdout
    ora = _oracle_run(po, scene, plan, images, rots)
    raw = np.fromfile(out_path, np.uint8)
    off, planes = _check_maps(raw, scene, plan, ora, po)
    assert off == raw.size
    # the worker wrote the reference's end-of-run dump (PM.cc:99-130, sigma <= 0.01)
    n_obj = sum(int(((s.astype(np.float64) <= 0.01) & (d > 0.000001)).sum()) for d, s in planes.values())
    with open(obj_path) as f:
        assert len(f.read().splitlines()) == n_obj


@pytest.mark.gpu
def test_class_reads_poses_at_call_time(tmp_path, pkg, po, synth, pipeline):
    """Key frames reach the device with a stale pose (uploaded before BA moved them); SetPose + one pass must give the
    result of the CURRENT poses, as the reference reads GetRotation()/GetTranslation() at call time
    (PM.cc:638-644, 854-860, 973-979)."""
    scene, plan, images, rots, _, scene_path = _host_scene(tmp_path, synth, pipeline, with_probes=False)
    out_path = tmp_path / "out.bin"
    r = subprocess.run([DRIVER, "--stale-pose", str(scene_path), str(out_path)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    ora = _oracle_run(po, scene, plan, images, rots)
    raw = np.fromfile(out_path, np.uint8)
    off, _ = _check_maps(raw, scene, plan, ora, po)
    assert off == raw.size


def _crafted_planes(rng, h, w, grow_case):
    depth = np.zeros((h, w), np.float32)
    sigma = np.zeros((h, w), np.float32)
    grad = rng.uniform(0, 20, (h, w)).astype(np.float32)
    valid = rng.random((h, w)) < 0.55
    base = 1.0 + 0.05 * np.sin(np.arange(w) / 9.0)[None, :] + 0.05 * np.cos(np.arange(h) / 7.0)[:, None]
    depth[valid] = (base + rng.normal(0, 0.01, (h, w)))[valid].astype(np.float32)
    sigma[valid] = rng.uniform(0.004, 0.03, (h, w))[valid].astype(np.float32)
    outl = valid & (rng.random((h, w)) < 0.1)
    depth[outl] *= np.float32(1.5)
    if grow_case:   # invalid pixels with a non-zero (and large) sigma: what the pipeline never produces
        #  -- ChiTest compares the neighbour's inverse depth with the centre's ~0, so only
        # sigmas of the order of the depth itself let anything through
        sigma[:] = rng.uniform(0.3, 2.0, (h, w)).astype(np.float32)
        depth[~valid & (rng.random((h, w)) < 0.5)] = np.float32(5e-7)
    return depth, sigma, grad


@pytest.mark.gpu
@pytest.mark.parametrize("h,w", [(48, 64), (37, 131), (240, 320)])
def test_intra_check_planes(ctx, po, h, w):
    depth, sigma, grad = _crafted_planes(np.random.default_rng(h * w), h, w, False)
    gd, gs = ctx.intra_planes(depth, sigma, grad, grow=False)
    od, os_ = po.intra_planes(depth, sigma)
    assert np.array_equal(gd, od) and np.array_equal(gs, os_)
    assert 0 < (od > 1e-6).sum() < (depth > 1e-6).sum()     # the check removed something and kept something


@pytest.mark.gpu
@pytest.mark.parametrize("h,w", [(48, 64), (37, 131), (240, 320)])
def test_intra_grow_planes(ctx, po, h, w):
    depth, sigma, grad = _crafted_planes(np.random.default_rng(h * w + 1), h, w, True)
    gd, gs = ctx.intra_planes(depth, sigma, grad, grow=True)
    od, os_ = po.intra_planes(depth, sigma, grad, grow=True)
    assert np.array_equal(gd, od, equal_nan=True) and np.array_equal(gs, os_, equal_nan=True)
    assert (od > 1e-6).sum() > (depth > 1e-6).sum()         # on these planes the stage does grow
    # and on pipeline-shaped planes (invalid => sigma 0) it is the identity
    depth2, sigma2, grad2 = _crafted_planes(np.random.default_rng(7), h, w, False)
    gd2, gs2 = ctx.intra_planes(depth2, sigma2, grad2, grow=True)
    assert np.array_equal(gd2, depth2) and np.array_equal(gs2, sigma2)


@pytest.mark.gpu
@pytest.mark.parametrize("window,threads", [(5, 3), (64, 1)])
def test_class_windowed_passes_equal_oracle(tmp_path, pkg, po, synth, pipeline, window, threads):
    """SemiDenseLoop issues a pass in windows (one sd_semidense_loop call each, asynchronous compact downloads, records
    expanded by host threads while the GPU runs the next window).  Three passes from scratch over a 24-key-frame map, in
    windows of 5 (key frames whose neighbours fall into a later window get their inter-key-frame check there) and in a
    single window: the KeyFrame members after the last pass are the oracle's, bit for bit (expansion over members that
    already hold the previous pass's values included).  This is the mode bench.py's `e2e_class` leg times."""
    import json
    scene, plan, images, rots, _, scene_path = _host_scene(tmp_path, synth, pipeline, nkf=24, n=5, seed=35, with_probes=False)
    out_path = tmp_path / "out.bin"
    r = subprocess.run([DRIVER, f"--bench=2,1,{window},{threads}", str(scene_path), str(out_path)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout + r.stderr
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("bench: ")][0]
    b = json.loads(line[len("bench: "):])
    assert b["keyframes"] == 24 and b["inter_checked"] == 24 and b["kf_per_s"] > 0
    ora = _oracle_run(po, scene, plan, images, rots)
    raw = np.fromfile(out_path, np.uint8)
    off, _ = _check_maps(raw, scene, plan, ora, po)
    assert off == raw.size
