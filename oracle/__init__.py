"""CPU oracle for the stereo hot path -- TEST INFRASTRUCTURE ONLY.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline / reference arm may
import this package.  The product (``multiview_b200``) never does.

Two layers live here:

* ``libmvoracle.so`` (``sgbm_oracle.c``, ``image_oracle.c``, ``geom_oracle.c``): the plain-C
  restatement of the reference algorithm, each function citing the reference file:line it follows.
* ``cv2ref``: thin calls into Python ``cv2`` (OpenCV 4.13.0 in this image), i.e. the same
  third-party dependency the reference links (pinned 4.4.0 there), used to PIN the restatement for
  a1-a4 (remap, grey, blur, StereoSGBM).  The fp64 geometry (a6-a10) has no runnable reference:
  parity unpinned (see geom_oracle.c header and DESIGN.md).
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libmvoracle.so")
_SRC = ["sgbm_oracle.c", "image_oracle.c", "geom_oracle.c", "census_oracle.c"]


def build(force: bool = False) -> str:
    """Compile the C oracle (gcc, -ffp-contract=off).  Returns the .so path."""
    srcs = [os.path.join(_HERE, s) for s in _SRC]
    stale = force or not os.path.exists(_SO) or any(os.path.getmtime(s) > os.path.getmtime(_SO) for s in srcs)
    if stale:
        subprocess.check_call(["make", "-s", "-B", "-C", _HERE, "libmvoracle.so"])
    return _SO


_lib44 = None


def lib44():
    """The oracle built with -DORC_SGBM_SUBTRACT_P2: the path recurrence of older OpenCV sources (the reference pins 4.4.0;
    the form cv2 4.13 computes, and that lib() restates, subtracts min(Lp) only).  Not pinned by any runnable OpenCV here."""
    global _lib44
    if _lib44 is None:
        so = os.path.join(_HERE, "libmvoracle44.so")
        srcs = [os.path.join(_HERE, s) for s in _SRC]
        if not os.path.exists(so) or any(os.path.getmtime(s) > os.path.getmtime(so) for s in srcs):
            subprocess.check_call(["make", "-s", "-B", "-C", _HERE, "libmvoracle44.so"])
        _lib44 = ctypes.CDLL(so)
    return _lib44


_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build())
        _lib.orc_point_cloud.restype = ctypes.c_int
    return _lib


class SgbmParams(ctypes.Structure):
    """Mirror of StereoSGBMDisparity::Params (stereosgbm_disp.hpp:23-33), OpenCV setter order."""

    _fields_ = [(n, ctypes.c_int) for n in ("block", "disp12_max_diff", "min_disp", "num_disp", "speckle_win",
                                            "speckle_range", "prefilter_cap", "uniqueness", "mode", "p1", "p2")]

    @classmethod
    def reference_defaults(cls, num_disp=96, mode=3):
        return cls(13, 596, 1, num_disp, 600, 250, 54, 3, mode, 22, 1000)


class Bcam(ctypes.Structure):
    _fields_ = [("krr_inv", ctypes.c_double * 9 * 2), ("c", ctypes.c_double * 3 * 2)]


class HistParams(ctypes.Structure):
    _fields_ = [("hist_sz", ctypes.c_float), ("min_z", ctypes.c_float), ("max_z", ctypes.c_float),
                ("weighting", ctypes.c_int), ("use_still", ctypes.c_int)]


def _p(a):
    return ctypes.c_void_p(a.ctypes.data) if a is not None else None


def _sz(v):
    return ctypes.c_size_t(int(v))


def sgbm_dims(p: SgbmParams, W: int):
    mn, w1 = ctypes.c_int(), ctypes.c_int()
    ok = lib().orc_sgbm_dims(ctypes.byref(p), W, ctypes.byref(mn), ctypes.byref(w1))
    return bool(ok), mn.value, w1.value


def sgbm_compute(p: SgbmParams, L: np.ndarray, R: np.ndarray, taps: bool = False, subtract_p2: bool = False):
    """StereoSGBM::compute restatement.  Returns disp16 (H,W) int16, or a dict with taps.  subtract_p2 selects the build
    with the older form of the path recurrence (lib44)."""
    L = np.ascontiguousarray(L, np.uint8)
    R = np.ascontiguousarray(R, np.uint8)
    H, W = L.shape
    ok, minx1, w1 = sgbm_dims(p, W)
    if not ok:
        raise ValueError("image too narrow for the disparity range")
    disp = np.empty((H, W), np.int16)
    C = S1 = S = raw = None
    if taps:
        C = np.empty((H, w1, p.num_disp), np.int16)
        S1 = np.empty_like(C)
        S = np.empty_like(C)
        raw = np.empty((H, W), np.int16)
    rc = (lib44() if subtract_p2 else lib()).orc_sgbm_compute(ctypes.byref(p), _p(L), _p(R), W, H, _sz(W), _p(disp), _p(C), _p(S1), _p(S), _p(raw))
    if rc:
        raise RuntimeError(f"orc_sgbm_compute rc={rc}")
    if taps:
        return {"disp": disp, "C": C, "S1": S1, "S": S, "raw": raw, "minX1": minx1, "W1": w1}
    return disp


def census_sgm(left: np.ndarray, right: np.ndarray, disparity_size: int = 256, taps: bool = False, vendored_edges: bool = False):
    """The reference's CUDA_SGM strategy (sgm_disp.cpp:102-211 over the vendored libSGM kernels), restated (see
    census_oracle.c).  vendored_edges: reproduce the vendored kernels' neighbour substitution at d = 0 / D-1 (64 and 128
    disparities; the form pinned against oracle/_ref/libsgm_ref.so).  left / right: uint8 or uint16 grey (H, W).  Returns
    the float disparity, or a dict with taps."""
    assert left.dtype == right.dtype and left.dtype in (np.uint8, np.uint16) and left.shape == right.shape
    left, right = np.ascontiguousarray(left), np.ascontiguousarray(right)
    H, W = left.shape
    out = np.empty((H, W), np.float32)
    S = np.empty((H, W, disparity_size), np.uint16) if taps else None
    d = np.empty((H, W), np.uint16) if taps else None
    rc = lib().orc_census_sgm_ex(_p(left), _p(right), 16 if left.dtype == np.uint16 else 8, W, H, disparity_size, _p(out), _p(S), _p(d),
                                 1 if vendored_edges else 0)
    if rc:
        raise RuntimeError(f"orc_census_sgm_ex rc={rc}")
    return {"disp": out, "S": S, "disp16": d} if taps else out


def disp_to_float(disp16: np.ndarray) -> np.ndarray:
    d = np.ascontiguousarray(disp16, np.int16)
    out = np.empty(d.shape, np.float32)
    lib().orc_disp_to_float(_p(d), d.shape[1], d.shape[0], _p(out))
    return out


def median3_s16(a):
    a = np.ascontiguousarray(a, np.int16)
    out = np.empty_like(a)
    H, W = a.shape
    lib().orc_median3_s16(_p(a), _sz(2 * W), W, H, _p(out), _sz(2 * W))
    return out


def filter_speckles(a, new_val, max_size, max_diff):
    a = np.array(a, np.int16, order="C")
    H, W = a.shape
    lib().orc_filter_speckles_s16(_p(a), _sz(2 * W), W, H, int(new_val), int(max_size), int(max_diff))
    return a


def remap_cubic_bgr8(src, mapx, mapy, border=(255, 255, 255)):
    src = np.ascontiguousarray(src, np.uint8)
    mapx = np.ascontiguousarray(mapx, np.float32)
    mapy = np.ascontiguousarray(mapy, np.float32)
    sh, sw, _ = src.shape
    h, w = mapx.shape
    dst = np.empty((h, w, 3), np.uint8)
    b = (ctypes.c_uint8 * 3)(*border)
    lib().orc_remap_cubic_bgr8(_p(src), sw, sh, _sz(sw * 3), _p(mapx), _p(mapy), _sz(w * 4), w, h, b, _p(dst), _sz(w * 3))
    return dst


def remap_bgr8(src, mapx, mapy, interp=2, border=(255, 255, 255)):
    """cv::remap with interp = cv::INTER_NEAREST 0 | LINEAR 1 | CUBIC 2 | LANCZOS4 4 (calc_rectified.hpp:17-22)."""
    src = np.ascontiguousarray(src, np.uint8)
    mapx = np.ascontiguousarray(mapx, np.float32)
    mapy = np.ascontiguousarray(mapy, np.float32)
    sh, sw, _ = src.shape
    h, w = mapx.shape
    dst = np.empty((h, w, 3), np.uint8)
    b = (ctypes.c_uint8 * 3)(*border)
    rc = lib().orc_remap_bgr8(int(interp), _p(src), sw, sh, _sz(sw * 3), _p(mapx), _p(mapy), _sz(w * 4), w, h, b, _p(dst), _sz(w * 3))
    if rc:
        raise ValueError(f"orc_remap_bgr8: unsupported interpolation {interp}")
    return dst


def bgr2gray(src):
    src = np.ascontiguousarray(src, np.uint8)
    h, w, _ = src.shape
    dst = np.empty((h, w), np.uint8)
    lib().orc_bgr2gray(_p(src), _sz(w * 3), w, h, _p(dst), _sz(w))
    return dst


def gauss_blur_u8(img, n=31, sigma=2.15):
    img = np.ascontiguousarray(img, np.uint8)
    h, w = img.shape
    dst = np.empty_like(img)
    rc = lib().orc_gauss_blur_u8(_p(img), _sz(w), w, h, int(n), ctypes.c_double(sigma), _p(dst), _sz(w))
    if rc:
        raise RuntimeError(f"orc_gauss_blur_u8 rc={rc}")
    return dst


def presmooth(img, sigma=2.15):
    """Disparity::calculate prologue, disparity.cpp:106-155: grey, then the 31x31 blur TWICE."""
    g = bgr2gray(img) if img.ndim == 3 else np.ascontiguousarray(img, np.uint8)
    if sigma > 0.0:
        for _ in range(2):
            g = gauss_blur_u8(g, 31, sigma)
    return g


def bcam_init(q, t, baseline, K) -> Bcam:
    out = Bcam()
    q = (ctypes.c_double * 4)(*q)
    t = (ctypes.c_double * 3)(*t)
    Kc = (ctypes.c_double * 9)(*np.asarray(K, np.float64).ravel())
    lib().orc_bcam_init(q, t, ctypes.c_double(baseline), Kc, ctypes.byref(out))
    return out


def solve3d(b: Bcam, x0, x1, y):
    out = (ctypes.c_double * 3)()
    lib().orc_solve3d(ctypes.byref(b), ctypes.c_double(x0), ctypes.c_double(x1), ctypes.c_double(y), out)
    return np.array(out[:])


def point_cloud(b: Bcam, disp: np.ndarray, left_is_ref=True):
    disp = np.ascontiguousarray(disp, np.float32)
    h, w = disp.shape
    Xs = np.empty((h * w, 3), np.float64)
    xy = np.empty((h * w, 2), np.int32)
    lookup = np.empty((h, w), np.int32)
    C = np.empty(3, np.float64)
    n = lib().orc_point_cloud(ctypes.byref(b), _p(disp), _sz(w * 4), w, h, int(left_is_ref), _p(Xs), _p(xy), _p(lookup), _p(C))
    return {"Xs": Xs[:n].copy(), "xy": xy[:n].copy(), "lookup": lookup, "C": C, "N": n}


def hist_dims(aabb, hist_sz):
    w, h = ctypes.c_int(), ctypes.c_int()
    a = (ctypes.c_double * 4)(*aabb)
    lib().orc_hist_dims(a, ctypes.c_float(hist_sz), ctypes.byref(w), ctypes.byref(h))
    return w.value, h.value


def floor_hist_camera(hp: HistParams, et, aabb, cam_center, Xs, xy, labels=None, still_score=None):
    """et = (tx,ty,tz, qx,qy,qz,qw, scale); aabb = (left, top, right, bottom)."""
    Xs = np.ascontiguousarray(Xs, np.float64)
    xy = np.ascontiguousarray(xy, np.int32)
    hw, hh = hist_dims(aabb, hp.hist_sz)
    hist = np.empty((hh, hw), np.float32)
    etc = (ctypes.c_double * 8)(*et)
    a = (ctypes.c_double * 4)(*aabb)
    cc = (ctypes.c_double * 3)(*cam_center)
    lw = lh = nl = 0
    if labels is not None:
        labels = np.ascontiguousarray(labels, np.int32)
        lh, lw = labels.shape
        still_score = np.ascontiguousarray(still_score, np.float64)
        nl = still_score.shape[0]
    lib().orc_floor_hist_camera(ctypes.byref(hp), etc, a, cc, _p(Xs), _p(xy), Xs.shape[0], _p(labels), lw, lh,
                                _p(still_score), nl, _p(hist), hw, hh)
    return hist


class SparseHistCell(ctypes.Structure):
    _fields_ = [("x", ctypes.c_int32), ("y", ctypes.c_int32), ("count", ctypes.c_float)]


def floor_hist_camera_cells(hp: HistParams, et, aabb, cam_center, Xs, xy, labels, still_score, super_pixel_min_z=0.25):
    """Histogram plus camera_info[cam].slic_to_hist_xy (floor-histogram.cpp:222-225,269-286) as an (n_labels, 3) float64
    array of (x, y, count)."""
    Xs = np.ascontiguousarray(Xs, np.float64)
    xy = np.ascontiguousarray(xy, np.int32)
    hw, hh = hist_dims(aabb, hp.hist_sz)
    hist = np.empty((hh, hw), np.float32)
    labels = np.ascontiguousarray(labels, np.int32)
    lh, lw = labels.shape
    still_score = np.ascontiguousarray(still_score, np.float64)
    nl = still_score.shape[0]
    cells = (SparseHistCell * nl)()
    f = lib().orc_floor_hist_camera_cells
    f.restype = None
    f(ctypes.byref(hp), (ctypes.c_double * 8)(*et), (ctypes.c_double * 4)(*aabb), (ctypes.c_double * 3)(*cam_center), _p(Xs), _p(xy),
      ctypes.c_int(Xs.shape[0]), _p(labels), ctypes.c_int(lw), ctypes.c_int(lh), _p(still_score), ctypes.c_int(nl), _p(hist),
      ctypes.c_int(hw), ctypes.c_int(hh), ctypes.c_float(super_pixel_min_z), cells)
    return hist, np.array([(c.x, c.y, c.count) for c in cells], np.float64).reshape(nl, 3)


def hist_sum(cam_hists):
    ch = np.ascontiguousarray(cam_hists, np.float32)
    n, hh, hw = ch.shape
    out = np.empty((hh, hw), np.float32)
    lib().orc_hist_sum(_p(ch), n, hw, hh, _p(out))
    return out


def disp_merge_ref(disp, ref, labels, still_cutoff):
    d = np.array(disp, np.float32, order="C")
    r = np.ascontiguousarray(ref, np.float32)
    lab = np.ascontiguousarray(labels, np.int32)
    cut = np.ascontiguousarray(still_cutoff, np.uint8)
    h, w = d.shape
    lib().orc_disp_merge_ref(_p(d), _p(r), _p(lab), _p(cut) if cut.size else None, int(cut.size), w, h)
    return d


def hist_post(hist, sigma=0.5, min_count=0.0, max_count=1000.0):
    """Tail of FloorHistogram::calculate (floor-histogram.cpp:411-434): blur + clip."""
    a = np.array(hist, np.float32, order="C")
    hh, hw = a.shape
    rc = lib().orc_hist_post(_p(a), hw, hh, ctypes.c_float(sigma), ctypes.c_float(min_count), ctypes.c_float(max_count))
    if rc:
        raise RuntimeError(f"orc_hist_post rc={rc}")
    return a


class LocParams(ctypes.Structure):
    """LocalizationData::Params fields used by the front half (localization-data.hpp:27-39)."""

    _fields_ = [(n, ctypes.c_double) for n in ("n_deviations", "person_diameter", "background_median", "background_stddev",
                                               "track_median", "track_stddev", "prob_bg_max")]

    @classmethod
    def reference_defaults(cls):
        return cls(3.0, 0.400, 0.0, 20.0, 200.0, 50.0, 0.050)


def localization_hists(p: LocParams, hist, hist_sz):
    """motion, radius-filtered, background, 3-D foreground and combined foreground histograms (localization-data.cpp:845-858)."""
    a = np.ascontiguousarray(hist, np.float32)
    hh, hw = a.shape
    out = [np.empty((hh, hw), np.float32) for _ in range(5)]
    lib().orc_localization_hists(ctypes.byref(p), _p(a), hw, hh, ctypes.c_double(hist_sz), *[_p(o) for o in out])
    return dict(zip(("motion", "rhist", "bg", "fg0", "fg"), out))


def input_images(raw, equalize=True):
    """input_images node (pipeline/nodes/input_images.cpp:65-76): (equal, grey)."""
    raw = np.ascontiguousarray(raw, np.uint8)
    h, w, _ = raw.shape
    equal = np.empty((h, w, 3), np.uint8)
    grey = np.empty((h, w), np.uint8)
    lib().orc_input_images(_p(raw), w, h, _sz(w * 3), int(bool(equalize)), _p(equal), _sz(w * 3), _p(grey), _sz(w))
    return equal, grey
