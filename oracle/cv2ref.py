"""cv2 (OpenCV 4.13.0) calls that pin the C restatement -- TEST INFRASTRUCTURE ONLY.

These are the same OpenCV entry points the reference calls on the hot path (reference pins 4.4.0,
built WITH_IPP=OFF; version skew is recorded in DESIGN.md):
  cv::remap        pipeline/nodes/calc_rectified.cpp:109-115
  cv::cvtColor     cost-functions/disparity/disparity.cpp:112
  cv::GaussianBlur cost-functions/disparity/disparity.cpp:141-152 (applied twice)
  cv::StereoSGBM   cost-functions/disparity/strategies/stereosgbm_disp.cpp:121-125
"""
from __future__ import annotations

import numpy as np


def _cv2():
    import cv2

    cv2.ipp.setUseIPP(False)
    return cv2


def remap_cubic_bgr8(src, mapx, mapy, border=(255, 255, 255)):
    cv2 = _cv2()
    return cv2.remap(src, mapx, mapy, cv2.INTER_CUBIC, borderMode=cv2.BORDER_CONSTANT, borderValue=tuple(border))


def remap_bgr8(src, mapx, mapy, interp, border=(255, 255, 255)):
    cv2 = _cv2()
    return cv2.remap(src, mapx, mapy, int(interp), borderMode=cv2.BORDER_CONSTANT, borderValue=tuple(border))


def presmooth(img, sigma=2.15):
    cv2 = _cv2()
    g = cv2.cvtColor(img, cv2.COLOR_BGR2GRAY) if img.ndim == 3 else img
    if sigma > 0.0:
        for _ in range(2):
            g = cv2.GaussianBlur(g, (31, 31), sigma, sigmaY=sigma, borderType=cv2.BORDER_DEFAULT)
    return g


def sgbm_compute(p, L, R):
    cv2 = _cv2()
    s = cv2.StereoSGBM_create(minDisparity=p.min_disp, numDisparities=p.num_disp, blockSize=p.block, P1=p.p1, P2=p.p2,
                              disp12MaxDiff=p.disp12_max_diff, preFilterCap=p.prefilter_cap,
                              uniquenessRatio=p.uniqueness, speckleWindowSize=p.speckle_win,
                              speckleRange=p.speckle_range, mode=p.mode)
    return s.compute(np.ascontiguousarray(L), np.ascontiguousarray(R))


def disp_to_float(d16):
    """stereosgbm_disp.cpp:141-150"""
    f = np.abs(d16.astype(np.float32)) / np.float32(16.0)
    f[f > np.float32(150.0)] = 0.0
    return f


def disparity_calculate(p, bgr0, bgr1, sigma=2.15):
    """Disparity::calculate with method STEREO_SGBM (disparity.cpp:68-182)."""
    return disp_to_float(sgbm_compute(p, presmooth(bgr0, sigma), presmooth(bgr1, sigma)))


def set_threads(n: int):
    _cv2().setNumThreads(int(n))


def hist_post(hist, sigma=0.5, min_count=0.0, max_count=1000.0):
    """floor-histogram.cpp:411-434 with the real cv::GaussianBlur."""
    cv2 = _cv2()
    import math

    a = np.array(hist, np.float32)
    if math.isfinite(sigma) and sigma > 0:
        n = int(math.ceil(sigma) * 4.0)
        n += 1 if n % 2 == 0 else 0
        a = cv2.GaussianBlur(a, (n, n), sigma, sigmaY=sigma, borderType=cv2.BORDER_DEFAULT)
    a[a < min_count] = 0.0
    a[a > max_count] = max_count
    return a


def input_images(raw, equalize=True):
    """hist_equalize (graphics/hist-equalize.cpp:9-30) + cvtColor BGR2GRAY, as input_images.cpp:65-76 calls them."""
    cv2 = _cv2()
    equal = raw
    if equalize:
        ycc = cv2.cvtColor(raw, cv2.COLOR_BGR2YCrCb)
        ch = list(cv2.split(ycc))
        ch[0] = cv2.equalizeHist(ch[0])
        equal = cv2.cvtColor(cv2.merge(ch), cv2.COLOR_YCrCb2BGR)
    return equal, cv2.cvtColor(equal, cv2.COLOR_BGR2GRAY)
