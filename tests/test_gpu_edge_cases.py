"""Edge cases of the disparity path: tiny and ragged images, strided host buffers, extreme parameters."""
import ctypes as C

import numpy as np
import pytest

import oracle
from multiview_b200 import api

pytestmark = pytest.mark.gpu


def _noise(h, w, seed):
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, (h, w), dtype=np.uint8), rng.integers(0, 256, (h, w), dtype=np.uint8)


def _smooth(h, w, seed):
    rng = np.random.default_rng(seed)
    base = rng.integers(0, 256, (h, w + 40), dtype=np.uint8).astype(np.float32)
    for _ in range(3):
        base = (base + np.roll(base, 1, 1) + np.roll(base, -1, 1) + np.roll(base, 1, 0)) / 4
    base = np.clip(base * 1.5 - 60, 0, 255).astype(np.uint8)
    return np.ascontiguousarray(base[:, 5:5 + w]), np.ascontiguousarray(base[:, :w])


def _check(ctx, L, R, **kw):
    D = kw.pop("D", 16)
    for mode in (1, 0, 3):
        sp = api.StereoSGBMParams(sg_num_disparities=D, sg_mode=mode, **kw)
        op = oracle.SgbmParams(sp.sg_SAD_window_size, sp.sg_12_max_diff, sp.sg_min_disparity, D, sp.sg_speckle_window_size,
                               sp.sg_speckle_range, sp.sg_prefilter_cap, sp.sg_uniqueness_ratio, mode, sp.sg_p1, sp.sg_p2)
        want = oracle.sgbm_compute(op, L, R)
        got = api.StereoSGBMDisparity.compute16(sp, L, R, ctx=ctx)
        assert np.array_equal(got, want), f"mode {mode} {L.shape}: {(got != want).sum()} px differ"


@pytest.mark.parametrize("h,w", [(1, 40), (2, 40), (3, 21), (5, 18), (17, 19), (33, 22), (16, 48), (31, 100)])
def test_tiny_and_ragged_images(ctx, h, w):
    """1..3 rows, 1..5 valid columns (W1 = w - 17), heights around the 16-row band size."""
    for gen in (_noise, _smooth):
        L, R = gen(h, w, h * 100 + w)
        _check(ctx, L, R, sg_SAD_window_size=5, sg_speckle_window_size=10, sg_speckle_range=2)


def test_disparity_range_wider_than_image_is_an_error(ctx):
    L, R = _noise(20, 30, 1)
    with pytest.raises(api.MvsError):
        api.StereoSGBMDisparity.compute16(api.StereoSGBMParams(sg_num_disparities=32), L, R, ctx=ctx)


def test_strided_host_buffers(ctx):
    """cv::Mat rows need not be packed: steps larger than the row are honoured on input and output."""
    L, R = _smooth(40, 120, 3)
    pad = 13
    Lp = np.full((40, 120 + pad), 7, np.uint8)
    Rp = np.full((40, 120 + pad), 9, np.uint8)
    Lp[:, :120], Rp[:, :120] = L, R
    out = np.full((40, 120 + 5), -999, np.int16)
    sp = api.StereoSGBMParams(sg_num_disparities=32, sg_mode=1, sg_SAD_window_size=7)
    cp = sp.to_c()
    rc = ctx.lib.mvs_sgbm_compute16(ctx.h, C.byref(cp), C.c_void_p(Lp.ctypes.data), C.c_void_p(Rp.ctypes.data), 120, 40, 120 + pad,
                                    C.c_void_p(out.ctypes.data), (120 + 5) * 2)
    ctx.check(rc, "mvs_sgbm_compute16")
    want = api.StereoSGBMDisparity.compute16(sp, L, R, ctx=ctx)
    assert np.array_equal(out[:, :120], want) and (out[:, 120:] == -999).all()


def test_extreme_parameters(ctx):
    L, R = _smooth(48, 160, 5)
    _check(ctx, L, R, D=64, sg_SAD_window_size=1, sg_p1=1, sg_p2=2, sg_uniqueness_ratio=0, sg_12_max_diff=1, sg_speckle_window_size=0)
    _check(ctx, L, R, D=32, sg_SAD_window_size=13, sg_p1=600, sg_p2=700, sg_uniqueness_ratio=99, sg_prefilter_cap=63,
           sg_speckle_window_size=5000, sg_speckle_range=1)
    _check(ctx, L, R, D=48, sg_SAD_window_size=3, sg_min_disparity=9, sg_uniqueness_ratio=50, sg_prefilter_cap=1)
    # all-equal images: every cost is zero, every disparity ties -> first minimum wins everywhere
    Z = np.full((24, 90), 128, np.uint8)
    _check(ctx, Z, Z.copy(), D=16, sg_SAD_window_size=3, sg_speckle_window_size=0)


def test_empty_and_degenerate_inputs(ctx):
    sp = api.StereoSGBMParams(sg_num_disparities=16)
    with pytest.raises(api.MvsError):
        api.Disparity.calculate(api.DisparityParams(stereosgbm_params=sp), True, np.zeros((0, 0), np.uint8), np.zeros((0, 0), np.uint8), ctx=ctx)
    with pytest.raises(api.MvsError):
        api.StereoSGBMDisparity.compute16(sp, np.zeros((4, 40), np.uint8), np.zeros((5, 40), np.uint8), ctx=ctx)
    # a cloud with no valid pixel and a histogram with no points
    bc = api.BinocularCamera().init((0, 0, 0, 1), (-1, 0, 0), 0.1, np.array([[100.0, 0, 8], [0, 100.0, 8], [0, 0, 1]]))
    pc = api.PointCloud().init(np.full((16, 16), np.nan, np.float32), True, bc, ctx=ctx)
    assert pc.N() == 0
    fh = api.FloorHistogram()
    api.FloorHistogram.calculate(api.FloorHistogramParams(weighting=api.WGT_NONE, apply_slic_still_weights=False), (-1, -1, 1, 1), 1,
                                 [(0, 0, 0, 0, 0, 0, 1, 1.0)], [pc], [(0, 0, 0)], fh, ctx=ctx)
    assert fh.hist.shape == (20, 20) and fh.hist.sum() == 0
