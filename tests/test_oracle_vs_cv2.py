"""Pins the oracle against the OpenCV build importable in this image (cv2), where available -- CPU only.

cv2 is the same third-party dependency the reference calls for a1-a4 (it pins 4.4.0; this image has 4.13.0).
On a machine without cv2 these tests skip and the committed golden vectors (test_oracle_golden.py) remain."""
import numpy as np
import pytest

import oracle
from multiview_b200 import synth

cv2 = pytest.importorskip("cv2")
from oracle import cv2ref  # noqa: E402


def _pair(W, H, D, seed, noise):
    if noise:
        rng = np.random.default_rng(seed)
        return rng.integers(0, 256, (H, W), dtype=np.uint8), rng.integers(0, 256, (H, W), dtype=np.uint8)
    l, r = synth.rectified_pair(W, H, D, seed)
    return oracle.bgr2gray(l), oracle.bgr2gray(r)


@pytest.mark.parametrize("mode", [0, 1, 3])
def test_sgbm_matrix(mode):
    cv2ref.set_threads(1)
    for (W, H, D) in [(72, 40, 16), (160, 56, 32)]:
        for (blk, p1, p2, ur, d12, sw, sr, cap, minD) in [(3, 8, 32, 0, 1, 0, 0, 54, 1), (9, 22, 1000, 3, 596, 0, 0, 54, 1),
                                                          (13, 22, 1000, 25, 1, 50, 2, 54, 1), (11, 22, 4000, 3, 596, 0, 0, 63, 1),
                                                          (7, 0, 0, -1, 0, 100, 1, 10, 0), (5, 10, 120, 5, 1, 20, 1, 31, 3)]:
            for noise in (False, True):
                L, R = _pair(W, H, D, W + mode + blk, noise)
                p = oracle.SgbmParams(blk, d12, minD, D, sw, sr, cap, ur, mode, p1, p2)
                assert np.array_equal(oracle.sgbm_compute(p, L, R), cv2ref.sgbm_compute(p, L, R)), (W, blk, noise)


def test_remap_and_presmooth():
    rng = np.random.default_rng(11)
    src = rng.integers(0, 256, (96, 128, 3), dtype=np.uint8)
    mx, my = synth.radial_maps(128, 96, 0.3, (2.0, 1.0))
    mx += rng.normal(0, 0.4, mx.shape).astype(np.float32)
    assert np.array_equal(oracle.remap_cubic_bgr8(src, mx, my), cv2ref.remap_cubic_bgr8(src, mx, my))
    for sigma in (2.15, 1.0, 3.3):
        assert np.array_equal(oracle.presmooth(src, sigma), cv2ref.presmooth(src, sigma))


def test_float_conversion():
    d = np.array([[-16, 0, 15, 2400, 2401, 32767, -32768]], np.int16)
    assert np.array_equal(oracle.disp_to_float(d), cv2ref.disp_to_float(d))


def test_hist_post_close_to_cv2():
    """Float blur: the restatement is within an ulp or two of cv2 (not bit-exact: OpenCV's float path depends on SIMD dispatch)."""
    rng = np.random.default_rng(3)
    h = ((rng.random((100, 120)) < 0.2) * rng.integers(1, 1500, (100, 120))).astype(np.float32)
    for sigma, mn, mx in [(0.5, 0.0, 1000.0), (1.3, 2.0, 500.0), (3.0, 0.0, 1e9)]:
        a, b = oracle.hist_post(h, sigma, mn, mx), cv2ref.hist_post(h, sigma, mn, mx)
        assert np.allclose(a, b, rtol=1e-6, atol=1e-4), np.abs(a - b).max()
    assert np.array_equal(oracle.hist_post(h, 0.0, 5.0, 700.0), cv2ref.hist_post(h, 0.0, 5.0, 700.0))


@pytest.mark.parametrize("interp", [0, 1, 2, 4])
def test_remap_all_interpolation_modes_match_cv2(interp):
    """calc_rectified.hpp:17-22: NEAREST / LINEAR / CUBIC / LANCZOS4, incl. border crossings and NaN / huge coordinates."""
    rng = np.random.default_rng(40 + interp)
    for (sw, sh, w, h) in [(80, 60, 70, 50), (33, 21, 64, 40)]:
        src = rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8)
        mx = (rng.random((h, w)) * (sw + 12) - 6).astype(np.float32)
        my = (rng.random((h, w)) * (sh + 12) - 6).astype(np.float32)
        mx[0, 0], my[0, 1], mx[1, 1], my[2, 2] = 1e9, -1e9, np.nan, 3e4
        mx[3, :8] = np.arange(8, dtype=np.float32) * 0.5  # exact halves: round-half-even
        a, b = oracle.remap_bgr8(src, mx, my, interp), cv2ref.remap_bgr8(src, mx, my, interp)
        assert np.array_equal(a, b), f"interp {interp}: {(a != b).any(axis=2).sum()} px differ"


def test_input_images_equalize_matches_cv2():
    """hist_equalize + BGR2GRAY of the input_images node, bit-exact with cv2."""
    rng = np.random.default_rng(31)
    imgs = [rng.integers(0, 256, (97, 131, 3), dtype=np.uint8), rng.integers(90, 140, (64, 80, 3), dtype=np.uint8),
            np.full((20, 24, 3), 77, np.uint8), synth.make_camera(320, 200, 64, config=3)["sensor"][0]]
    imgs[0][:30, :40] = rng.integers(0, 30, (30, 40, 3))
    for img in imgs:
        for eq in (True, False):
            a, b = oracle.input_images(img, eq), cv2ref.input_images(img, eq)
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
