"""Round-2 entry points of the C ABI against the oracle and against the entry points they shortcut:

  * sd_interkf_check_points  == sd_interkf_check + sd_update_pointset           (PM.cc:302-304 in one launch)
  * sd_semidense_loop        == the per-key-frame calls                          (PM.cc:137-307 in one C call)
  * sd_download_compact      == sd_download, bit for bit, after expansion        (PM.cc:225-226, 361-363 leave dense planes)
  * SdParams.interkf_inplace == the oracle's in-place mode                       (PM.cc:262, 764, 793)
  * re-upload + sd_import_maps with work still queued (race on the halo pattern)
  * BASELINE configs 3 and 5 at FULL size (720p N=20; 1080p N=10/20/40 with the widened prior), bitwise
"""
import ctypes as C

import numpy as np
import pytest

from helpers import SMALL_CAM, run_gpu, run_oracle

pytestmark = pytest.mark.gpu


def _setup(synth, pipeline, cam, nkf, n, seed=1234, **kw):
    scene = synth.Scene(cam, seed=seed)
    plan = pipeline.SequencePlan(scene, nkf, n, **kw)
    images = {k: scene.render(k) for k in plan.ids}
    return scene, plan, images


def test_fused_check_points_equals_two_calls(pkg, po, synth, pipeline):
    scene, plan, images = _setup(synth, pipeline, SMALL_CAM, 9, 6, seed=41)
    a = run_gpu(pkg, pipeline, plan, images)                      # sd_interkf_check + sd_update_pointset
    b = run_gpu(pkg, pipeline, plan, images, interkf=False)
    for k in plan.ids:
        b.interkf_check_points(k, plan.nbrs[k])
    b.sync()
    ora = run_oracle(po, plan, images)
    for k in plan.ids:
        for which in (po.DEPTH_CHECKED, po.SIGMA, po.POINTS):
            assert np.array_equal(a.plane(k, which), b.plane(k, which)), (k, which)
            assert np.array_equal(b.plane(k, which), ora.plane(k, which)), (k, which)
    # a second fused check of the same key frame starts again from the intra map and from a zeroed point plane
    b.interkf_check_points(plan.ids[4], plan.nbrs[plan.ids[4]])
    assert np.array_equal(b.plane(plan.ids[4], po.POINTS), ora.plane(plan.ids[4], po.POINTS))
    assert (b.plane(plan.ids[4], po.DEPTH_CHECKED) > 1e-6).sum() > 100
    a.close(); b.close()


@pytest.mark.parametrize("with_points", [False, True])
def test_compact_download_equals_dense(pkg, po, synth, pipeline, with_points):
    scene, plan, images = _setup(synth, pipeline, "480p", 8, 6, seed=43)
    ctx = run_gpu(pkg, pipeline, plan, images)
    for k in plan.ids:
        rec, n = ctx.download_compact(k, points=with_points)
        assert n == len(rec) and 0 < n < (scene.w - 4) * (scene.h - 4)
        assert len(np.unique(rec["pixel"])) == n                              # one record per pixel
        dense = ctx.download(k, points=with_points)
        got = pkg.expand_compact(rec, scene.h, scene.w, points=with_points)
        for g, d in zip(got, dense):
            assert np.array_equal(g, d)
        # every pixel the inter-KF check zeroed keeps its sigma (PM.cc:764) and is therefore still listed
        assert ((rec["depth"] == 0) & (rec["sigma"] != 0)).sum() == ((dense[0] == 0) & (dense[1] != 0)).sum()
    # capacity smaller than the list: count reports what is available, the copy stops at capacity
    rec, n = ctx.download_compact(plan.ids[3], points=with_points, capacity=100)
    assert len(rec) == 100 and n > 100
    # before the inter-KF stage the records carry the intra-KF map
    ctx2 = run_gpu(pkg, pipeline, plan, images, interkf=False)
    rec, n = ctx2.download_compact(plan.ids[2])
    d, s = pkg.expand_compact(rec, scene.h, scene.w)
    assert np.array_equal(d, ctx2.plane(plan.ids[2], po.DEPTH)) and np.array_equal(s, ctx2.plane(plan.ids[2], po.SIGMA))
    ctx.close(); ctx2.close()


def test_compact_download_async_ring_and_loop(pkg, po, synth, pipeline):
    """More pending compact downloads than the staging ring holds (16), issued through sd_semidense_loop; the
    results land in pinned memory at sd_sync and equal the step-by-step path and the oracle."""
    import torch
    scene, plan, images = _setup(synth, pipeline, SMALL_CAM, 28, 5, seed=47)
    ora = run_oracle(po, plan, images)
    cap = (scene.w - 4) * (scene.h - 4)
    ctx = pkg.Context(0)
    pipeline.upload_all(ctx, plan, images)
    recs = torch.zeros((len(plan.ids), cap * pkg.COMPACT_POINT.itemsize), dtype=torch.uint8).pin_memory()
    counts = torch.full((len(plan.ids),), -1, dtype=torch.int32).pin_memory()
    items = [dict(kf_id=k, nbrs=plan.nbrs[k], rots=plan.rots[k], prior=plan.prior[k],
                  download=(pkg.SD_DL_COMPACT_POINTS, recs[i].data_ptr(), cap, counts[i:].data_ptr()))
             for i, k in enumerate(plan.ids)]
    loop = ctx.make_loop(items)
    for rep in range(2):                                   # the second pass re-runs on resident planes with downloads pending
        done = ctx.semidense_loop(loop)
        ctx.sync()
        assert done.all()
        for i, k in enumerate(plan.ids):
            n = int(counts[i])
            assert 0 < n <= cap
            rec = recs[i].numpy().view(pkg.COMPACT_POINT)[:n]
            d, s, p = pkg.expand_compact(rec, scene.h, scene.w, points=True)
            assert np.array_equal(d, ora.plane(k, po.DEPTH_CHECKED)), (rep, k)
            assert np.array_equal(s, ora.plane(k, po.SIGMA)), (rep, k)
            assert np.array_equal(p, ora.plane(k, po.POINTS)), (rep, k)
    ctx.close()


def test_back_to_back_passes_without_sync(pkg, po, synth, pipeline):
    """bench.py's step, three times with NO host synchronisation in between and different images every time: upload of
    all key frames (the planes of the previous pass are still being read), sd_semidense_loop with compact downloads.
    Every cross-stream dependency is a single wait on the latest record needed (upload / maps-ready / K4 numbers):
    a missing one shows up as planes of the wrong pass.  The last pass must equal the oracle on ITS images."""
    import torch
    nkf, n = 14, 5
    scenes = [synth.Scene(SMALL_CAM, seed=s) for s in (61, 62, 63)]
    plans = [pipeline.SequencePlan(sc, nkf, n) for sc in scenes]
    cap = (scenes[0].w - 4) * (scenes[0].h - 4)
    ctx = pkg.Context(0)
    host = [torch.from_numpy(np.stack([sc.render(k) for k in pl.ids])).pin_memory() for sc, pl in zip(scenes, plans)]
    recs = torch.zeros((3, nkf, cap * pkg.COMPACT_POINT.itemsize), dtype=torch.uint8).pin_memory()
    counts = torch.full((3, nkf), -1, dtype=torch.int32).pin_memory()
    loops = [ctx.make_loop([dict(kf_id=k, nbrs=pl.nbrs[k], rots=pl.rots[k], prior=pl.prior[k],
                                 download=(pkg.SD_DL_COMPACT_POINTS, recs[r, i].data_ptr(), cap, counts[r, i:].data_ptr()))
                            for i, k in enumerate(pl.ids)]) for r, pl in enumerate(plans)]
    for r, (sc, pl) in enumerate(zip(scenes, plans)):
        for i, k in enumerate(pl.ids):
            ctx.upload_keyframe_ptr(k, host[r][i].data_ptr(), sc.w, sc.h, sc.w, sc.K, pl.poses[k], device=False)
        assert ctx.semidense_loop(loops[r]).all()
    ctx.sync()
    for r, (sc, pl) in enumerate(zip(scenes, plans)):
        images = {k: host[r][i].numpy() for i, k in enumerate(pl.ids)}
        ora = run_oracle(po, pl, images)
        for i, k in enumerate(pl.ids):
            m = int(counts[r, i])
            assert 0 < m <= cap
            d, s, p = pkg.expand_compact(recs[r, i].numpy().view(pkg.COMPACT_POINT)[:m], sc.h, sc.w, points=True)
            assert np.array_equal(d, ora.plane(k, po.DEPTH_CHECKED)), (r, k)
            assert np.array_equal(s, ora.plane(k, po.SIGMA)), (r, k)
            assert np.array_equal(p, ora.plane(k, po.POINTS)), (r, k)
        if r == 2:   # the resident planes are the last pass's
            for k in pl.ids:
                assert np.array_equal(ctx.plane(k, po.DEPTH), ora.plane(k, po.DEPTH)), k
    # the same key frame searched twice in a row without an upload in between (k1_search puts its task counter back itself)
    k = plans[2].ids[6]
    for _ in range(2):
        ctx.semidense(k, plans[2].nbrs[k], plans[2].rots[k], *plans[2].prior[k])
    assert np.array_equal(ctx.plane(k, po.DEPTH), ora.plane(k, po.DEPTH))
    assert np.array_equal(ctx.plane(k, po.SIGMA), ora.plane(k, po.SIGMA))
    ctx.close()


def test_semidense_loop_dense_downloads_and_skips(pkg, po, synth, pipeline):
    """sd_semidense_loop with dense downloads; a list that lacks some neighbours skips their inter-KF stage like
    PM.cc:289-297 and says so."""
    import torch
    scene, plan, images = _setup(synth, pipeline, SMALL_CAM, 10, 4, seed=53)
    ora = run_oracle(po, plan, images)
    ctx = pkg.Context(0)
    pipeline.upload_all(ctx, plan, images)
    host = torch.zeros((len(plan.ids), 5, scene.h, scene.w), dtype=torch.float32).pin_memory()
    items = [dict(kf_id=k, nbrs=plan.nbrs[k], rots=plan.rots[k], prior=plan.prior[k],
                  download=(pkg.SD_DL_DENSE_POINTS, host[i, 0].data_ptr(), host[i, 1].data_ptr(), host[i, 2].data_ptr()))
             for i, k in enumerate(plan.ids)]
    done = ctx.semidense_loop(ctx.make_loop(items))
    ctx.sync()
    assert done.all()
    for i, k in enumerate(plan.ids):
        assert np.array_equal(host[i, 0].numpy(), ora.plane(k, po.DEPTH_CHECKED))
        assert np.array_equal(host[i, 1].numpy(), ora.plane(k, po.SIGMA))
        assert np.array_equal(host[i, 2:5].numpy().reshape(-1)[:3 * scene.h * scene.w].reshape(scene.h, 3 * scene.w),
                              ora.plane(k, po.POINTS))
    # only the first half of the list: key frames whose neighbours are in the second half have no maps to check against
    ctx2 = pkg.Context(0)
    pipeline.upload_all(ctx2, plan, images)
    half = [dict(kf_id=k, nbrs=plan.nbrs[k], rots=plan.rots[k], prior=plan.prior[k]) for k in plan.ids[:5]]
    done = ctx2.semidense_loop(ctx2.make_loop(half))
    ctx2.sync()
    expect = [all(int(j) in plan.ids[:5] for j in plan.nbrs[k]) for k in plan.ids[:5]]
    assert list(done.astype(bool)) == expect and not all(expect) and any(expect)
    for k, e in zip(plan.ids[:5], expect):
        assert np.array_equal(ctx2.plane(k, po.DEPTH), ora.plane(k, po.DEPTH))
        if e:
            assert np.array_equal(ctx2.plane(k, po.DEPTH_CHECKED), ora.plane(k, po.DEPTH_CHECKED))
    # a skipped item that asked for compact records WITH points has no point set: its count comes back 0, not stale
    ctx3 = pkg.Context(0)
    pipeline.upload_all(ctx3, plan, images)
    cap = (scene.w - 4) * (scene.h - 4)
    recs = torch.zeros((5, cap * pkg.COMPACT_POINT.itemsize), dtype=torch.uint8).pin_memory()
    counts = torch.full((5,), 12345, dtype=torch.int32).pin_memory()
    half = [dict(kf_id=k, nbrs=plan.nbrs[k], rots=plan.rots[k], prior=plan.prior[k],
                 download=(pkg.SD_DL_COMPACT_POINTS, recs[i].data_ptr(), cap, counts[i:].data_ptr()))
            for i, k in enumerate(plan.ids[:5])]
    done = ctx3.semidense_loop(ctx3.make_loop(half))
    ctx3.sync()
    assert list(done.astype(bool)) == expect
    for i, (k, e) in enumerate(zip(plan.ids[:5], expect)):
        if not e:
            assert int(counts[i]) == 0
        else:
            m = int(counts[i])
            assert 0 < m <= cap
            d, _, p = pkg.expand_compact(recs[i].numpy().view(pkg.COMPACT_POINT)[:m], scene.h, scene.w, points=True)
            assert np.array_equal(d, ora.plane(k, po.DEPTH_CHECKED)) and np.array_equal(p, ora.plane(k, po.POINTS))
    ctx.close(); ctx2.close(); ctx3.close()


def test_interkf_inplace_mode(pkg, po, synth, pipeline):
    """PM.cc:764, 793 write depth_map_ in place inside the loop over key frames (PM.cc:262-307): a key frame visited later
    reads neighbours that were already smoothed / zeroed.  Ascending visiting order on both sides."""
    scene, plan, images = _setup(synth, pipeline, SMALL_CAM, 12, 6, seed=59)
    gp, op = pkg.default_params(), po.default_params()
    gp.interkf_inplace = 1
    op.interkf_inplace = 1
    ora = run_oracle(po, plan, images, params=op)
    snap = run_oracle(po, plan, images)
    ctx = run_gpu(pkg, pipeline, plan, images, params=gp)
    differs = 0
    for k in plan.ids:
        for which in (po.DEPTH, po.SIGMA, po.DEPTH_CHECKED, po.POINTS):
            assert np.array_equal(ctx.plane(k, which), ora.plane(k, which)), (k, which)
        differs += int(np.count_nonzero(ora.plane(k, po.DEPTH_CHECKED) != snap.plane(k, po.DEPTH_CHECKED)))
    assert differs > 0                                      # the mode is not a no-op on this scene
    # the fused entry obeys the mode as well
    ctx2 = run_gpu(pkg, pipeline, plan, images, params=gp, interkf=False)
    for k in plan.ids:
        ctx2.interkf_check_points(k, plan.nbrs[k])
    ctx2.sync()
    for k in plan.ids:
        assert np.array_equal(ctx2.plane(k, po.DEPTH_CHECKED), ora.plane(k, po.DEPTH_CHECKED))
    # and so does sd_semidense_loop: in this mode it issues the checks in list order after all the searches (PM.cc:262's
    # order), not as soon as an item's neighbours are queued
    ctx3 = pkg.Context(0, gp)
    pipeline.upload_all(ctx3, plan, images)
    items = [dict(kf_id=k, nbrs=plan.nbrs[k], rots=plan.rots[k], prior=plan.prior[k]) for k in plan.ids]
    assert ctx3.semidense_loop(ctx3.make_loop(items)).all()
    ctx3.sync()
    for k in plan.ids:
        for which in (po.DEPTH_CHECKED, po.POINTS):
            assert np.array_equal(ctx3.plane(k, which), ora.plane(k, which)), (k, which)
    ctx.close(); ctx2.close(); ctx3.close()


def test_reupload_then_import_with_work_queued(pkg, po, synth, pipeline):
    """The halo pattern of a multi-GPU step, repeated without sd_sync: re-upload a key frame (which forgets the slot's
    reader bookkeeping), then install its maps, while the inter-KF checks of the previous step that READ those maps are
    still queued.  Every step must reproduce the oracle."""
    import torch
    scene, plan, images = _setup(synth, pipeline, "480p", 9, 6, seed=61)
    ora = run_oracle(po, plan, images)
    src = run_gpu(pkg, pipeline, plan, images, interkf=False)   # donor of the maps ("the peer rank")
    halo = plan.ids[:3]
    bufs = {k: torch.empty((2, scene.h, scene.w), dtype=torch.float32, device="cuda:0") for k in halo}
    for k in halo:
        src.export_maps(k, bufs[k][0].data_ptr(), bufs[k][1].data_ptr())
    src.sync()
    garbage = torch.full((2, scene.h, scene.w), 7.0, dtype=torch.float32, device="cuda:0")
    ctx = pkg.Context(0)
    pipeline.upload_all(ctx, plan, images)
    own = [k for k in plan.ids if k not in halo]
    for step in range(4):
        for k in halo:                                       # step N+1 starts while step N's checks are queued
            ctx.upload_keyframe(k, images[k], scene.K, plan.poses[k])
            if step % 2:                                     # a wrong intermediate import must not leak into earlier readers
                ctx.import_maps(k, garbage[0].data_ptr(), garbage[1].data_ptr())
            ctx.import_maps(k, bufs[k][0].data_ptr(), bufs[k][1].data_ptr())
        for k in own:
            ctx.semidense(k, plan.nbrs[k], plan.rots[k], *plan.prior[k])
        for k in own:
            ctx.interkf_check_points(k, plan.nbrs[k])
    ctx.sync()
    for k in own:
        assert np.array_equal(ctx.plane(k, po.DEPTH_CHECKED), ora.plane(k, po.DEPTH_CHECKED)), k
        assert np.array_equal(ctx.plane(k, po.POINTS), ora.plane(k, po.POINTS)), k
    src.close(); ctx.close()


def test_import_maps_host_and_confident_fetch(pkg, po, synth, pipeline):
    scene, plan, images = _setup(synth, pipeline, SMALL_CAM, 8, 6, seed=67)
    ora = run_oracle(po, plan, images)
    ctx = run_gpu(pkg, pipeline, plan, images)
    k = plan.ids[3]
    d, s = ctx.download(k)
    # release, come back from the host copies (what ProbabilityMapping does for a key frame that left the device)
    ctx.release_keyframe(k)
    ctx.upload_keyframe(k, images[k], scene.K, plan.poses[k])
    ctx.import_maps_host(k, d, s)
    ctx.update_pointset(k)
    assert np.array_equal(ctx.plane(k, po.POINTS), ora.plane(k, po.POINTS))
    with pytest.raises(pkg.SdError):                          # imported maps have no compact form
        ctx.download_compact(k)
    # device-resident selection, then one fetch
    px, ps, pp, n, nconf = ctx.confident_points_device(plan.ids[2], 0.05, 0)
    xyz = np.empty((n, 3), np.float32)
    got = C.c_int()
    ctx._ck(ctx.L.sd_confident_points_fetch(ctx.h, n, xyz.ctypes.data, None, None, C.byref(got)))
    ox = ora.confident_points(plan.ids[2], 0.05, 0)[0]
    assert got.value == n == len(ox) and np.array_equal(xyz, ox)
    ctx.close()


# ---- BASELINE configs 3 and 5 at full size -------------------------------------------------------------------------
def _partial_run(pkg, po, plan, images, sd_ids, chk_ids):
    """K0 for every key frame, K1..K3 for `sd_ids`, K4+K5 for `chk_ids`, on both sides."""
    scene = plan.scene
    ctx = pkg.Context(0)
    ora = po.Oracle(threads=po.lib().ora_max_threads())
    for k in plan.ids:
        ctx.upload_keyframe(k, images[k], scene.K, plan.poses[k])
        ora.add_keyframe(k, images[k], scene.K, plan.poses[k])
    for k in sd_ids:
        ctx.semidense(k, plan.nbrs[k], plan.rots[k], *plan.prior[k])
        ora.semidense(k, plan.nbrs[k], plan.rots[k], *plan.prior[k])
    for k in chk_ids:
        ctx.interkf_check_points(k, plan.nbrs[k])
        ora.interkf(k, plan.nbrs[k])
        ora.pointset(k)
    ctx.sync()
    return ctx, ora


def _bitwise(ctx, ora, po, k, planes):
    for which in planes:
        g, o = ctx.plane(k, which), ora.plane(k, which)
        assert np.array_equal(g, o), f"kf {k} plane {which}: {np.count_nonzero(g != o)} differing values"


def test_config3_full_size_720p_n20(pkg, po, synth, pipeline):
    """BASELINE.json configs[2] geometry: 1280x720, 20 covisible neighbours (k1_fuse_rows).  GradImg, GradTheta, depth and
    sigma of the key frames around the middle bitwise; the middle key frame also through the inter-KF check and the
    point set (its 20 neighbours all carry maps)."""
    n = 20
    scene, plan, images = _setup(synth, pipeline, "720p", 2 * n + 1, n, seed=71)
    mid = plan.ids[n]
    sd_ids = sorted({mid, *[int(j) for j in plan.nbrs[mid]]})
    ctx, ora = _partial_run(pkg, po, plan, images, sd_ids, [mid])
    for k in (mid, sd_ids[0]):
        _bitwise(ctx, ora, po, k, (po.GRAD, po.THETA, po.DEPTH, po.SIGMA))
    _bitwise(ctx, ora, po, mid, (po.DEPTH_CHECKED, po.POINTS))
    assert (ctx.plane(mid, po.DEPTH_CHECKED) > 1e-6).mean() > 0.05
    ctx.close()


@pytest.mark.parametrize("n", [10, 20, 40])
def test_config5_full_size_1080p_wide_prior(pkg, po, synth, pipeline, n):
    """BASELINE.json configs[4]: 1920x1080, N in {10, 20, 40}, inverse-depth prior widened to mean +- 4 sigma.
    Two key frames through K0..K3, bitwise (N=5 at this size is test_gpu_parity.test_1080p_n5)."""
    scene, plan, images = _setup(synth, pipeline, "1080p", n + 1, n, seed=73 + n, prior_sigmas=4.0)
    a, b = plan.ids[n // 2], plan.ids[0]
    ctx, ora = _partial_run(pkg, po, plan, images, [a, b], [])
    for k in (a, b):
        _bitwise(ctx, ora, po, k, (po.GRAD, po.THETA, po.DEPTH, po.SIGMA))
        assert (ctx.plane(k, po.DEPTH) > 1e-6).mean() > 0.03
    ctx.close()
