"""sd_intra_depth_check_planes_static: the one C-ABI entry point that allocates no device memory.

Some hosts of this library may allocate on the GPU only while they start up: a tracking / local-mapping thread with a
frame deadline cannot sit through cudaMalloc and cudaFree (the latter waits for everything the GPU is doing, including the
key-frame pipeline in flight), and a process that gave all device memory to its own pool at start-up, or runs under a
memory quota, may not get any.  For them the library carries the working planes of the intra-key-frame checking stage
inside its own device image (SD_STATIC_MAX_PIXELS pixels, include/sdpm.h) and this entry point runs the stage in them.

What is checked here, through the C ABI: the static variant returns exactly (bit for bit, like every other parity test of
this repository) what sd_intra_depth_check_planes returns; a second call right after the first -- on different planes of a
different size, so the shared array still holds the first call's data -- is just as exact, and so is a repeat of the first;
and planes larger than the built-in array are refused with a message naming the limit, leaving the caller's planes
untouched.  The file is named so that pytest collects it last (a failure here cannot hide the other tests under -x), and it
makes no assumption about what ran before it in the process: it passes alone and after the whole suite.
"""
import numpy as np
import pytest


def _planes(seed, h, w):
    """Depth / sigma / gradient planes with ~55 % valid pixels and some outliers, so that the stage both keeps and drops."""
    rng = np.random.default_rng(seed)
    depth = np.zeros((h, w), np.float32)
    sigma = np.zeros((h, w), np.float32)
    grad = rng.uniform(0, 20, (h, w)).astype(np.float32)
    valid = rng.random((h, w)) < 0.55
    base = 1.0 + 0.05 * np.sin(np.arange(w) / 9.0)[None, :] + 0.05 * np.cos(np.arange(h) / 7.0)[:, None]
    depth[valid] = (base + rng.normal(0, 0.01, (h, w)))[valid].astype(np.float32)
    sigma[valid] = rng.uniform(0.004, 0.03, (h, w))[valid].astype(np.float32)
    outliers = valid & (rng.random((h, w)) < 0.1)
    depth[outliers] *= np.float32(1.5)
    return depth, sigma, grad


def _same_bits(a, b):
    return a.shape == b.shape and np.array_equal(a.view(np.uint32), b.view(np.uint32))


@pytest.mark.gpu
def test_static_variant_matches_allocating_entry_point_twice_in_a_row(ctx, pkg):
    first = _planes(11, 240, 320)     # float4 store path (w % 4 == 0)
    second = _planes(12, 37, 131)     # odd size: scalar stores, and smaller than what the array now holds
    want_first = ctx.intra_planes(*first, grow=False)     # sd_intra_depth_check_planes (cudaMalloc per call)
    want_second = ctx.intra_planes(*second, grow=False)
    assert 0 < (want_first[0] > 1e-6).sum() < (first[0] > 1e-6).sum()   # the stage did something on these planes

    got_first = ctx.intra_check_planes_static(*first)
    got_second = ctx.intra_check_planes_static(*second)   # straight after: the array still holds call one's planes
    got_again = ctx.intra_check_planes_static(*first)
    for got, want in ((got_first, want_first), (got_second, want_second), (got_again, want_first)):
        assert _same_bits(got[0], want[0]) and _same_bits(got[1], want[1])


@pytest.mark.gpu
def test_static_variant_at_its_largest_size_and_from_a_second_context(ctx, pkg):
    h, w = 720, 1280
    assert h * w == pkg.SD_STATIC_MAX_PIXELS
    planes = _planes(13, h, w)
    want = ctx.intra_planes(*planes, grow=False)
    got = ctx.intra_check_planes_static(*planes)
    assert _same_bits(got[0], want[0]) and _same_bits(got[1], want[1])
    other = pkg.Context(0)            # another context on the same GPU shares the array (calls are sequential here)
    try:
        got2 = other.intra_check_planes_static(*planes)
    finally:
        other.close()
    assert _same_bits(got2[0], want[0]) and _same_bits(got2[1], want[1])


@pytest.mark.gpu
def test_static_variant_refuses_planes_beyond_its_limit(ctx, pkg):
    h, w = 721, 1280
    depth, sigma, grad = _planes(14, h, w)
    with pytest.raises(pkg.SdError) as err:
        ctx.intra_check_planes_static(depth, sigma, grad)
    assert err.value.code == -1      # SD_ERR_INVALID_ARG
    assert "SD_STATIC_MAX_PIXELS" in str(err.value) and str(pkg.SD_STATIC_MAX_PIXELS) in str(err.value)
    # the allocating entry point still takes them
    d, s = ctx.intra_planes(depth, sigma, grad, grow=False)
    assert d.shape == (h, w) and np.isfinite(d).all()
