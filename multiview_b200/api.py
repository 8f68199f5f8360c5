"""Host-side mirror of the reference's operator interface for the stereo hot path, over the C ABI.

Names, argument meaning and error behaviour follow the reference (paths under
multiview/multiview_cpp/src/perceive/):

  calc_rectified()                    pipeline/nodes/calc_rectified.cpp:66-132
  Disparity.calculate()               cost-functions/disparity/disparity.hpp:44-57, disparity.cpp:68-182
  StereoSGBMDisparity.calculate()     cost-functions/disparity/strategies/stereosgbm_disp.hpp:36-41
  BinocularCamera.init / solve3d      geometry/projective/binocular-camera.cpp:682-785,908-927
  PointCloud.init()                   geometry/point-cloud.hpp:31-33
  FloorHistogram.calculate()          cost-functions/floor-histogram.hpp:117-131

The reference's functions are noexcept and signal failure by logging and leaving the output untouched
(stereosgbm_disp.cpp:96-99,132-138); here failures raise MvsError carrying mvs_last_error().
All compute happens in libmvstereo.so on the GPU; numpy arrays are only the host containers
(cv::Mat / std::vector stand-ins).
"""
from __future__ import annotations

import ctypes as C
import math
from dataclasses import dataclass, field

import numpy as np

from . import _lib
from ._lib import Bcam, FrameDesc, FrameResult, HistParams, LocParams, MvsError, SgbmParams

INTER_NEAREST, INTER_LINEAR, INTER_CUBIC, INTER_LANCZOS4 = 0, 1, 2, 4
MODE_SGBM, MODE_HH, MODE_SGBM_3WAY, MODE_HH4 = 0, 1, 2, 3
WGT_NONE, WGT_NORM, WGT_QUADRANCE = 0, 1, 2


def _vp(a):
    return C.c_void_p(a.ctypes.data) if a is not None else None


class Context:
    """One CUDA stream + device arena (mvs_ctx).  Use one per host thread / camera."""

    def __init__(self, device: int = 0):
        self.lib = _lib.load()
        h = C.c_void_p()
        rc = self.lib.mvs_ctx_create(device, C.byref(h))
        if rc:
            raise MvsError(f"mvs_ctx_create(device={device}) failed with {rc}: no usable CUDA device (no CPU fallback)")
        self.h = h
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.lib.mvs_ctx_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def check(self, rc, what):
        if rc:
            raise MvsError(f"{what} failed ({rc}): {self.lib.mvs_last_error(self.h).decode()}")

    def sync(self):
        self.check(self.lib.mvs_sync(self.h), "mvs_sync")

    def launch_count(self) -> int:
        return int(self.lib.mvs_launch_count(self.h))

    def stage_ms(self, stage: str) -> float:
        return float(self.lib.mvs_stage_ms(self.h, stage.encode()))


_default_ctx = None


def default_context() -> Context:
    global _default_ctx
    if _default_ctx is None:
        _default_ctx = Context(0)
    return _default_ctx


# ---------------------------------------------------------------------------------------------- a1
def calc_rectified(image, mapx, mapy, interpolation_method=INTER_CUBIC, ctx: Context | None = None):
    """cv::remap(in, rectified, mapx, mapy, method, BORDER_CONSTANT, Scalar(255,255,255))."""
    ctx = ctx or default_context()
    image = np.ascontiguousarray(image, np.uint8)
    mapx = np.ascontiguousarray(mapx, np.float32)
    mapy = np.ascontiguousarray(mapy, np.float32)
    if image.ndim != 3 or image.shape[2] != 3 or mapx.shape != mapy.shape:
        raise MvsError("calc_rectified: expected a BGR image and two maps of equal size")
    sh, sw, _ = image.shape
    h, w = mapx.shape
    dst = np.empty((h, w, 3), np.uint8)
    white = (C.c_uint8 * 3)(255, 255, 255)
    ctx.check(ctx.lib.mvs_remap_bgr8(ctx.h, _vp(image), sw, sh, sw * 3, _vp(mapx), _vp(mapy), w * 4, w, h,
                                     int(interpolation_method), white, _vp(dst), w * 3), "mvs_remap_bgr8")
    return dst


# ---------------------------------------------------------------------------------------------- a2-a4
@dataclass
class StereoSGBMParams:
    """StereoSGBMDisparity::Params, strategies/stereosgbm_disp.hpp:23-33 (same names and defaults)."""
    sg_SAD_window_size: int = 13
    sg_12_max_diff: int = 596
    sg_min_disparity: int = 1
    sg_num_disparities: int = 96
    sg_speckle_window_size: int = 600
    sg_speckle_range: int = 250
    sg_prefilter_cap: int = 54
    sg_uniqueness_ratio: int = 3
    sg_mode: int = 3
    sg_p1: int = 22
    sg_p2: int = 1000

    def to_c(self) -> SgbmParams:
        # set_ssgbm_params (stereosgbm_disp.cpp:55-80): unknown modes fall back to MODE_SGBM
        mode = self.sg_mode if self.sg_mode in (MODE_SGBM, MODE_HH, MODE_SGBM_3WAY, MODE_HH4) else MODE_SGBM
        return SgbmParams(self.sg_SAD_window_size, self.sg_12_max_diff, self.sg_min_disparity, self.sg_num_disparities,
                          self.sg_speckle_window_size, self.sg_speckle_range, self.sg_prefilter_cap,
                          self.sg_uniqueness_ratio, mode, self.sg_p1, self.sg_p2)


class StereoSGBMDisparity:
    Params = StereoSGBMParams

    @staticmethod
    def compute16(p: StereoSGBMParams, grey0, grey1, ctx: Context | None = None):
        """cv::StereoSGBM::compute -> CV_16S Q4 disparity."""
        ctx = ctx or default_context()
        g0 = np.ascontiguousarray(grey0, np.uint8)
        g1 = np.ascontiguousarray(grey1, np.uint8)
        if g0.ndim != 2 or g0.shape != g1.shape:
            raise MvsError("expected a pair of grey images")
        h, w = g0.shape
        out = np.empty((h, w), np.int16)
        cp = p.to_c()
        ctx.check(ctx.lib.mvs_sgbm_compute16(ctx.h, C.byref(cp), _vp(g0), _vp(g1), w, h, w, _vp(out), w * 2), "mvs_sgbm_compute16")
        return out

    @staticmethod
    def calculate(p: StereoSGBMParams, l_is_ref_image: bool, grey0, grey1, ctx: Context | None = None):
        """-> CV_32FC1 disparity (|d16|/16, > 150 -> 0), stereosgbm_disp.cpp:84-156."""
        if not l_is_ref_image:
            raise MvsError("right-reference matching (ximgproc::createRightMatcher) is never used by the pipeline "
                           "(disparity.cpp:171-174 passes true) and is not implemented")
        return Disparity._run(p, 0.0, grey0, grey1, ctx)

    @staticmethod
    def volumes(p: StereoSGBMParams, grey0, grey1, ctx: Context | None = None):
        """Test tap: (C, S, raw16) of the last stages, unpadded [h][w1][D]."""
        ctx = ctx or default_context()
        g0 = np.ascontiguousarray(grey0, np.uint8)
        g1 = np.ascontiguousarray(grey1, np.uint8)
        h, w = g0.shape
        w1 = w - p.sg_min_disparity - p.sg_num_disparities
        Cv = np.empty((h, w1, p.sg_num_disparities), np.int16)
        Sv = np.empty_like(Cv)
        raw = np.empty((h, w), np.int16)
        cp = p.to_c()
        ctx.check(ctx.lib.mvs_sgbm_volumes(ctx.h, C.byref(cp), _vp(g0), _vp(g1), w, h, w, _vp(Cv), _vp(Sv), _vp(raw)), "mvs_sgbm_volumes")
        return Cv, Sv, raw


STEREO_BM, STEREO_SGBM, TRICLOPS, CUDA_SGM = 0, 1, 2, 3


@dataclass
class DisparityParams:
    """Disparity::Params, disparity.hpp:26-41 (only the members the SGBM path reads)."""
    disparity_method: int = STEREO_SGBM
    use_cuda_if_available: bool = True
    presmooth_sigma: float = 2.15
    stereosgbm_params: StereoSGBMParams = field(default_factory=StereoSGBMParams)


class Disparity:
    Params = DisparityParams

    @staticmethod
    def _run(sp: StereoSGBMParams, sigma: float, im0, im1, ctx):
        ctx = ctx or default_context()
        a = np.ascontiguousarray(im0, np.uint8)
        b = np.ascontiguousarray(im1, np.uint8)
        if a.size == 0 or b.size == 0:
            raise MvsError("attempt to calculate disparity on empty image")
        if a.shape != b.shape or a.ndim not in (2, 3) or (a.ndim == 3 and a.shape[2] != 3):
            raise MvsError("expected a pair of grey or BGR images of equal size")
        ch = 1 if a.ndim == 2 else 3
        h, w = a.shape[:2]
        out = np.empty((h, w), np.float32)
        cp = sp.to_c()
        ctx.check(ctx.lib.mvs_disparity_sgbm(ctx.h, C.byref(cp), float(sigma), _vp(a), _vp(b), ch, w, h, w * ch, _vp(out), w * 4),
                  "mvs_disparity_sgbm")
        return out

    @staticmethod
    def calculate(p: DisparityParams, l_is_ref_image: bool, image0, image1, method: int | None = None, ctx: Context | None = None):
        """Disparity::calculate: grey conversion, double 31x31 blur, then the selected strategy -> float image."""
        method = p.disparity_method if method is None else method
        if method != STEREO_SGBM:
            raise MvsError("only STEREO_SGBM is on the accelerated path (STEREO_BM / TRICLOPS / CUDA_SGM are out of scope)")
        if not l_is_ref_image:
            raise MvsError("l_is_ref_image=false is never used by the pipeline and is not implemented")
        return Disparity._run(p.stereosgbm_params, p.presmooth_sigma, image0, image1, ctx)


class SGMDisparity:
    """strategies/sgm_disp.hpp:13-35: the CUDA_SGM strategy (libSGM: census + Hamming semi-global matching)."""

    @staticmethod
    def calculate(disparity_size: int, l_is_ref_image: bool, image0, image1, ctx: Context | None = None, tap: bool = False):
        if not l_is_ref_image:
            raise MvsError("SGMDisparity: the pipeline only calls this with the left image as reference")
        ctx = ctx or default_context()
        image0, image1 = np.ascontiguousarray(image0), np.ascontiguousarray(image1)
        if image0.dtype != image1.dtype or image0.dtype not in (np.uint8, np.uint16) or image0.ndim != 2 or image0.shape != image1.shape:
            raise MvsError("expected a pair of 8-bit or 16-bit greyscale images of equal size")   # sgm_disp.cpp:122-132
        h, w = image0.shape
        bpp = image0.dtype.itemsize
        out = np.empty((h, w), np.float32)
        d16 = np.empty((h, w), np.uint16) if tap else None
        ctx.check(ctx.lib.mvs_disparity_census(ctx.h, _vp(image0), _vp(image1), 8 * bpp, w, h, w * bpp, int(disparity_size), _vp(out), w * 4,
                                               _vp(d16)), "mvs_disparity_census")
        return (out, d16) if tap else out


def presmooth(image, sigma=2.15, ctx: Context | None = None):
    """The prologue of Disparity::calculate alone (grey + double blur), for stage-level tests."""
    ctx = ctx or default_context()
    a = np.ascontiguousarray(image, np.uint8)
    ch = 1 if a.ndim == 2 else 3
    h, w = a.shape[:2]
    out = np.empty((h, w), np.uint8)
    ctx.check(ctx.lib.mvs_presmooth(ctx.h, _vp(a), ch, w, h, w * ch, float(sigma), _vp(out), w), "mvs_presmooth")
    return out


def disp_map_update(disparity, ref_disparity, slic_lookup, still_cuttoff, ctx: Context | None = None):
    """The merge of disp_map_update::Task::execute (pipeline/nodes/disp_map_update.cpp:98-113): returns the
    disparity with holes / non-still super-pixels filled from the scene's reference disparity."""
    ctx = ctx or default_context()
    d = np.array(disparity, np.float32, order="C")
    r = np.ascontiguousarray(ref_disparity, np.float32)
    lab = np.ascontiguousarray(slic_lookup, np.int32)
    cut = np.ascontiguousarray(still_cuttoff, np.uint8)
    if d.shape != r.shape or d.shape != lab.shape or d.ndim != 2:
        raise MvsError("disp_map_update: disparity, reference disparity and SLIC lookup must have the same size")
    h, w = d.shape
    ctx.check(ctx.lib.mvs_disp_merge_ref(ctx.h, _vp(d), w * 4, _vp(r), w * 4, _vp(lab), w * 4, _vp(cut) if cut.size else None,
                                         int(cut.size), w, h), "mvs_disp_merge_ref")
    return d


# ---------------------------------------------------------------------------------------------- a6-a7
class BinocularCamera:
    """The part of BinocularCamera the path needs: init() -> KRR_inv_[2], C_[2] (host side, numpy),
    binocular-camera.cpp:682-785 with use_calib_roi=false."""

    def __init__(self):
        self.KRR_inv = [np.zeros((3, 3)), np.zeros((3, 3))]
        self.C = [np.zeros(3), np.zeros(3)]

    def init(self, q, t, baseline, K):
        x, y, z, w = q
        R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                      [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                      [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]], np.float64)
        Rinv = np.linalg.inv(R)
        t = np.asarray(t, np.float64)
        tn = t / np.linalg.norm(t)
        C0 = np.zeros(3)
        C1 = -(Rinv @ (tn * baseline))
        z0 = R[2, :]
        X = (C1 - C0) / np.linalg.norm(C1 - C0)
        Y = np.cross(z0, X)
        Y /= np.linalg.norm(Y)
        Z = np.cross(X, Y)
        Z /= np.linalg.norm(Z)
        Rn = np.stack([X, Y, Z])
        K = np.asarray(K, np.float64)
        KRR0 = K @ Rn
        KRR1 = K @ (Rn @ Rinv) @ R
        self.KRR_inv = [np.linalg.inv(KRR0), np.linalg.inv(KRR1)]
        self.C = [C0, C1]
        return self

    def to_c(self) -> Bcam:
        b = Bcam()
        for i in range(2):
            for k, v in enumerate(np.asarray(self.KRR_inv[i], np.float64).ravel()):
                b.krr_inv[i][k] = v
            for k in range(3):
                b.c[i][k] = float(self.C[i][k])
        return b

    @classmethod
    def from_c(cls, b: Bcam):
        o = cls()
        o.KRR_inv = [np.array(b.krr_inv[i][:]).reshape(3, 3) for i in range(2)]
        o.C = [np.array(b.c[i][:]) for i in range(2)]
        return o


class PointCloud:
    """PointCloud (geometry/point-cloud.hpp:19-36): Xs, xy, lookup, C."""

    def __init__(self):
        self.Xs = np.zeros((0, 3), np.float64)
        self.xy = np.zeros((0, 2), np.int32)
        self.lookup = np.zeros((0, 0), np.int32)
        self.C = np.zeros(3, np.float64)

    def N(self):
        return self.Xs.shape[0]

    def init(self, disparity, left_is_reference: bool, bino_cam, ctx: Context | None = None):
        ctx = ctx or default_context()
        d = np.ascontiguousarray(disparity, np.float32)
        if d.ndim != 2:
            raise MvsError("PointCloud.init: expected a CV_32FC1 disparity image")
        h, w = d.shape
        Xs = np.empty((h * w, 3), np.float64)
        xy = np.empty((h * w, 2), np.int32)
        self.lookup = np.empty((h, w), np.int32)
        n = C.c_int32(0)
        b = bino_cam.to_c() if hasattr(bino_cam, "to_c") else bino_cam
        ctx.check(ctx.lib.mvs_point_cloud(ctx.h, C.byref(b), _vp(d), w * 4, w, h, int(bool(left_is_reference)), _vp(Xs), _vp(xy),
                                          _vp(self.lookup), C.byref(n), _vp(self.C)), "mvs_point_cloud")
        self.Xs = Xs[:n.value].copy()
        self.xy = xy[:n.value].copy()
        return self


# ---------------------------------------------------------------------------------------------- a9-a10
@dataclass
class FloorHistogramParams:
    """FloorHistogram::Params, cost-functions/floor-histogram.hpp:23-60 (members used by the hot loop)."""
    hist_sz: float = 0.10
    hist_min_z: float = 0.50
    hist_max_z: float = 2.00
    min_hist_count: float = 0.0
    max_hist_count: float = 1000.0
    gaussian_sigma: float = 0.50
    weighting: int = WGT_NORM
    apply_slic_still_weights: bool = True
    skip_cameras: tuple = ()
    super_pixel_min_z: float = 0.25
    ordered: bool = False   # not a reference field: sum every bin in point order (weighted histograms bit-exact)

    def to_c(self) -> HistParams:
        return HistParams(self.hist_sz, self.hist_min_z, self.hist_max_z, int(self.weighting), int(self.apply_slic_still_weights),
                          int(self.ordered), self.super_pixel_min_z)


def hist_dims(aabb, hist_sz):
    lib = _lib.load()
    w, h = C.c_int(), C.c_int()
    a = (C.c_double * 4)(*aabb)
    if lib.mvs_hist_dims(a, hist_sz, C.byref(w), C.byref(h)):
        raise MvsError("bad histogram bounds")
    return w.value, h.value


class FloorHistogram:
    """FloorHistogram::calculate (cost-functions/floor-histogram.cpp:105-300, 411-434): per-camera histograms, their
    sum in camera order, Gaussian smoothing and clipping.  `hist_raw` keeps the sum before smoothing."""

    Params = FloorHistogramParams

    def __init__(self):
        self.bounds = None
        self.hist_sz = 0.05
        self.hist = None
        self.hist_raw = None
        self.camera_hists = []

    @staticmethod
    def calculate(p: FloorHistogramParams, aabb, n_cams, transforms, point_clouds, translations, floor_hist,
                  slic_lookups=None, still_scores=None, ctx: Context | None = None) -> bool:
        """aabb = (left, top, right, bottom); transforms[i] = (tx,ty,tz,qx,qy,qz,qw,scale)."""
        ctx = ctx or default_context()
        if not all(math.isfinite(v) for v in aabb) or (aabb[2] - aabb[0]) * (aabb[3] - aabb[1]) <= 0.0:
            raise MvsError("Cannot calculate floor histogram, because scene bounding box does not have finite area")
        hw, hh = hist_dims(aabb, p.hist_sz)
        cp = p.to_c()
        a = (C.c_double * 4)(*aabb)
        hists = np.zeros((n_cams, hh, hw), np.float32)
        for cam in range(n_cams):
            if cam in p.skip_cameras:
                continue
            pc = point_clouds[cam]
            et = (C.c_double * 8)(*transforms[cam])
            cc = (C.c_double * 3)(*translations[cam])
            Xs = np.ascontiguousarray(pc.Xs, np.float64)
            xy = np.ascontiguousarray(pc.xy, np.int32)
            lab = still = None
            lw = lh = nl = 0
            if slic_lookups is not None and slic_lookups[cam] is not None and still_scores is not None:
                lab = np.ascontiguousarray(slic_lookups[cam], np.int32)
                lh, lw = lab.shape
                still = np.ascontiguousarray(still_scores[cam], np.float64)
                nl = still.shape[0]
            ctx.check(ctx.lib.mvs_floor_hist_camera(ctx.h, C.byref(cp), et, a, cc, _vp(Xs), _vp(xy), Xs.shape[0], _vp(lab), lw, lh,
                                                    _vp(still), nl, _vp(hists[cam]), hw, hh), "mvs_floor_hist_camera")
        total = np.empty((hh, hw), np.float32)
        ctx.check(ctx.lib.mvs_hist_sum(ctx.h, _vp(hists), n_cams, hw, hh, _vp(total)), "mvs_hist_sum")
        floor_hist.bounds = tuple(aabb)
        floor_hist.hist_sz = float(np.float32(p.hist_sz))
        floor_hist.hist_raw = total.copy()
        ctx.check(ctx.lib.mvs_hist_post(ctx.h, _vp(total), hw, hh, float(p.gaussian_sigma), float(p.min_hist_count),
                                        float(p.max_hist_count)), "mvs_hist_post")
        floor_hist.hist = total
        floor_hist.camera_hists = [hists[i] for i in range(n_cams)]
        return True


# ---------------------------------------------------------------------------------------------- frame path
def make_frame_desc(cam: dict, sgbm: StereoSGBMParams, hist: FloorHistogramParams, aabb, sigma=2.15, on_device=False,
                    dev_ptrs=None, interpolation_method=INTER_CUBIC, left_is_reference=True, slic_lookup=None,
                    still_score=None) -> FrameDesc:
    """Build a mvs_frame_desc from a synth.make_camera() dict (host pointers) or device pointers.  slic_lookup (int32 h x w)
    and still_score (float64 per label) are the inputs of the reference's default histogram weighting; the caller keeps
    them (and the camera's arrays) alive while the descriptor is in use."""
    d = FrameDesc()
    d.interp, d.left_is_ref = int(interpolation_method), int(bool(left_is_reference))
    if slic_lookup is not None and still_score is not None:
        assert slic_lookup.dtype == np.int32 and slic_lookup.flags.c_contiguous and still_score.dtype == np.float64
        d.labels, d.lstep = slic_lookup.ctypes.data, slic_lookup.strides[0]
        d.still_score, d.n_labels, d.labels_on_device = still_score.ctypes.data, int(still_score.size), 0
    w, h = cam["W"], cam["H"]
    if on_device:
        for i in range(2):
            d.sensor[i], d.mapx[i], d.mapy[i] = dev_ptrs["sensor"][i], dev_ptrs["mapx"][i], dev_ptrs["mapy"][i]
    else:
        for i in range(2):
            d.sensor[i] = cam["sensor"][i].ctypes.data
            d.mapx[i] = cam["mapx"][i].ctypes.data
            d.mapy[i] = cam["mapy"][i].ctypes.data
    d.sw, d.sh, d.w, d.h = w, h, w, h
    d.sstep, d.mstep = w * 3, w * 4
    d.inputs_on_device = int(on_device)
    d.maps_on_device = int(on_device)
    d.sgbm = sgbm.to_c()
    d.presmooth_sigma = sigma
    d.bcam = BinocularCamera().init(cam["q"], cam["t"], cam["baseline"], cam["K"]).to_c()
    d.hist = hist.to_c()
    for k in range(8):
        d.et[k] = cam["et"][k]
    for k in range(4):
        d.aabb[k] = aabb[k]
    for k in range(3):
        d.cam_center[k] = cam["et"][k]
    return d


def maps_register(ctx: Context, mapx, mapy) -> int:
    """mvs_maps_register: the per-video conversion of a sensor's rectification maps (get_xy_mapping's result) into the
    fixed-point form the cubic remap reads; returns the handle for FrameDesc.maps_handle[i]."""
    mapx, mapy = np.ascontiguousarray(mapx, np.float32), np.ascontiguousarray(mapy, np.float32)
    h, w = mapx.shape
    handle = C.c_int(0)
    ctx.check(ctx.lib.mvs_maps_register(ctx.h, _vp(mapx), _vp(mapy), w * 4, w, h, 0, C.byref(handle)), "mvs_maps_register")
    return handle.value


def frame_run(ctx: Context, slot: int, desc: FrameDesc) -> FrameResult:
    res = FrameResult()
    ctx.check(ctx.lib.mvs_frame_run(ctx.h, slot, C.byref(desc), C.byref(res)), "mvs_frame_run")
    return res


def frame_run_batch(ctx: Context, first_slot: int, descs) -> list:
    """The camera-frames of one step together (mvs_frame_run_batch); descs is a sequence of FrameDesc."""
    n = len(descs)
    arr = descs if isinstance(descs, C.Array) else (FrameDesc * n)(*descs)
    res = (FrameResult * n)()
    ctx.check(ctx.lib.mvs_frame_run_batch(ctx.h, first_slot, arr, n, res), "mvs_frame_run_batch")
    return list(res)


def frame_fetch(ctx: Context, slot: int, w: int, h: int, hw: int, hh: int, want_cloud=True):
    disp = np.empty((h, w), np.float32)
    lookup = np.empty((h, w), np.int32)
    Xs = np.empty((h * w, 3), np.float64) if want_cloud else None
    xy = np.empty((h * w, 2), np.int32) if want_cloud else None
    cen = np.empty(3, np.float64)
    hist = np.empty((hh, hw), np.float32)
    n = C.c_int32(0)
    ctx.check(ctx.lib.mvs_frame_fetch(ctx.h, slot, _vp(disp), w * 4, _vp(Xs), _vp(xy), _vp(lookup), C.byref(n), _vp(cen), _vp(hist)),
              "mvs_frame_fetch")
    out = {"disp": disp, "lookup": lookup, "N": n.value, "C": cen, "cam_hist": hist}
    if want_cloud:
        out["Xs"], out["xy"] = Xs[:n.value], xy[:n.value]
    return out


def frame_fetch_batch(ctx: Context, first_slot: int, n: int, w: int, h: int, hw: int, hh: int, want_cloud=True, n_labels=0):
    """mvs_frame_fetch_batch into fresh numpy arrays: one blocking point per batch (two with clouds)."""
    from ._lib import FetchDesc

    outs, descs, ns = [], (FetchDesc * n)(), [C.c_int32(0) for _ in range(n)]
    for i in range(n):
        o = {"disp": np.empty((h, w), np.float32), "lookup": np.empty((h, w), np.int32), "C": np.empty(3, np.float64),
             "cam_hist": np.empty((hh, hw), np.float32)}
        if want_cloud:
            o["Xs"], o["xy"] = np.empty((h * w, 3), np.float64), np.empty((h * w, 2), np.int32)
        descs[i].disp, descs[i].dstep, descs[i].lookup = o["disp"].ctypes.data, w * 4, o["lookup"].ctypes.data
        descs[i].centroid, descs[i].cam_hist = o["C"].ctypes.data, o["cam_hist"].ctypes.data
        descs[i].n_out = C.pointer(ns[i])
        if want_cloud:
            descs[i].Xs, descs[i].xy = o["Xs"].ctypes.data, o["xy"].ctypes.data
        if n_labels:
            o["_cells"] = (_lib.SparseHistCell * n_labels)()
            descs[i].slic_to_hist_xy = C.cast(o["_cells"], C.c_void_p)
        outs.append(o)
    ctx.check(ctx.lib.mvs_frame_fetch_batch(ctx.h, first_slot, n, descs), "mvs_frame_fetch_batch")
    ctx.sync()
    for i, o in enumerate(outs):
        o["N"] = ns[i].value
        if want_cloud:
            o["Xs"], o["xy"] = o["Xs"][:o["N"]], o["xy"][:o["N"]]
        if n_labels:   # (n_labels, 3) float64: x, y, count of every SparseHistCell
            o["slic_to_hist_xy"] = np.array([(c.x, c.y, c.count) for c in o.pop("_cells")], np.float64).reshape(n_labels, 3)
    return outs


# ---------------------------------------------------------------------------------------- localization (front half)
@dataclass
class LocalizationParams:
    """LocalizationData::Params, the fields its front half reads (localization-data.hpp:27-39, same names and defaults)."""
    use_median: bool = False      # ignored by the reference too (localization-data.cpp:213-219)
    n_deviations: float = 3.0
    person_diameter: float = 0.400
    background_median: float = 0.0
    background_stddev: float = 20.0
    track_median: float = 200.0
    track_stddev: float = 50.0
    prob_bg_max: float = 0.050

    def to_c(self) -> LocParams:
        return LocParams(self.n_deviations, self.person_diameter, self.background_median, self.background_stddev,
                         self.track_median, self.track_stddev, self.prob_bg_max)


class LocalizationHists:
    """What LocalizationData::calculate derives from the floor histogram before anything else
    (localization-data.cpp:845-850, pose-less branch :855-858): motion_hist, the radius-filtered histogram, bg, fg0, fg."""

    @staticmethod
    def calculate(p: LocalizationParams, hist, hist_sz: float, ctx: Context | None = None):
        ctx = ctx or default_context()
        a = np.ascontiguousarray(hist, np.float32)
        hh, hw = a.shape
        out = {k: np.empty((hh, hw), np.float32) for k in ("motion", "rhist", "bg", "fg0", "fg")}
        pc = p.to_c()
        ctx.check(ctx.lib.mvs_localization_hists(ctx.h, C.byref(pc), _vp(a), hw, hh, float(hist_sz), *[_vp(v) for v in out.values()]),
                  "mvs_localization_hists")
        return out


# ------------------------------------------------------------------------------------------------- input images
def input_images(raw, raw_image_equalize: bool = False, ctx: Context | None = None):
    """input_images::Task::execute (pipeline/nodes/input_images.cpp:65-76): returns (equal, grey)."""
    ctx = ctx or default_context()
    raw = np.ascontiguousarray(raw, np.uint8)
    if raw.ndim != 3 or raw.shape[2] != 3:
        raise ValueError("input_images: BGR 8UC3 image expected")
    h, w, _ = raw.shape
    equal = np.empty((h, w, 3), np.uint8)
    grey = np.empty((h, w), np.uint8)
    ctx.check(ctx.lib.mvs_input_images(ctx.h, _vp(raw), w, h, w * 3, int(bool(raw_image_equalize)), _vp(equal), w * 3, _vp(grey), w),
              "mvs_input_images")
    return equal, grey
