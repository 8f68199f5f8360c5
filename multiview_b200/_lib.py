"""ctypes binding of libmvstereo.so (the C ABI declared in include/mvstereo.h).

There is no CPU fallback: if the shared library is missing or no CUDA device is present, every entry
point raises.  Build with ``python -m multiview_b200.build`` (nvcc, sm_100a).
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("MVS_LIB") or os.path.join(_HERE, "libmvstereo.so")   # MVS_LIB: a developer build of the same library

c_int, c_size_t, c_void_p, c_double, c_float = C.c_int, C.c_size_t, C.c_void_p, C.c_double, C.c_float


class SgbmParams(C.Structure):
    _fields_ = [(n, c_int) for n in ("block", "disp12_max_diff", "min_disp", "num_disp", "speckle_win", "speckle_range",
                                     "prefilter_cap", "uniqueness", "mode", "p1", "p2")]


class Bcam(C.Structure):
    _fields_ = [("krr_inv", c_double * 9 * 2), ("c", c_double * 3 * 2)]


class HistParams(C.Structure):
    _fields_ = [("hist_sz", c_float), ("min_z", c_float), ("max_z", c_float), ("weighting", c_int), ("use_still", c_int),
                ("ordered", c_int), ("super_pixel_min_z", c_float)]


class SparseHistCell(C.Structure):
    _fields_ = [("x", C.c_int32), ("y", C.c_int32), ("count", c_float)]


class LocParams(C.Structure):
    _fields_ = [(n, c_double) for n in ("n_deviations", "person_diameter", "background_median", "background_stddev",
                                        "track_median", "track_stddev", "prob_bg_max")]


class FrameDesc(C.Structure):
    _fields_ = [("sensor", c_void_p * 2), ("mapx", c_void_p * 2), ("mapy", c_void_p * 2),
                ("sw", c_int), ("sh", c_int), ("w", c_int), ("h", c_int),
                ("sstep", c_size_t), ("mstep", c_size_t), ("inputs_on_device", c_int), ("maps_on_device", c_int),
                ("sgbm", SgbmParams), ("presmooth_sigma", c_double), ("bcam", Bcam), ("hist", HistParams),
                ("et", c_double * 8), ("aabb", c_double * 4), ("cam_center", c_double * 3),
                ("interp", c_int), ("left_is_ref", c_int), ("labels", c_void_p), ("lstep", c_size_t), ("still_score", c_void_p),
                ("n_labels", c_int), ("labels_on_device", c_int), ("maps_handle", c_int * 2)]


class FetchDesc(C.Structure):
    _fields_ = [("disp", c_void_p), ("dstep", c_size_t), ("Xs", c_void_p), ("xy", c_void_p), ("lookup", c_void_p),
                ("n_out", C.POINTER(C.c_int32)), ("centroid", c_void_p), ("cam_hist", c_void_p), ("slic_to_hist_xy", c_void_p)]


class FrameResult(C.Structure):
    _fields_ = [("rectified_dev", c_void_p * 2), ("disp16_dev", c_void_p), ("disp_dev", c_void_p), ("Xs_dev", c_void_p),
                ("xy_dev", c_void_p), ("lookup_dev", c_void_p), ("cam_hist_dev", c_void_p), ("n_dev", c_void_p),
                ("centroid_dev", c_void_p), ("hw", c_int), ("hh", c_int), ("slic_to_hist_xy_dev", c_void_p), ("n_labels", c_int)]


# name -> (restype, argtypes); must list every symbol include/mvstereo.h declares
SIGNATURES = {
    "mvs_ctx_create": (c_int, [c_int, C.POINTER(c_void_p)]),
    "mvs_ctx_destroy": (None, [c_void_p]),
    "mvs_last_error": (C.c_char_p, [c_void_p]),
    "mvs_sync": (c_int, [c_void_p]),
    "mvs_stream": (c_void_p, [c_void_p]),
    "mvs_tail_stream": (c_void_p, [c_void_p]),
    "mvs_version": (C.c_char_p, []),
    "mvs_launch_count": (C.c_longlong, [c_void_p]),
    "mvs_remap_bgr8": (c_int, [c_void_p, c_void_p, c_int, c_int, c_size_t, c_void_p, c_void_p, c_size_t, c_int, c_int,
                               c_int, c_void_p, c_void_p, c_size_t]),
    "mvs_maps_register": (c_int, [c_void_p, c_void_p, c_void_p, c_size_t, c_int, c_int, c_int, C.POINTER(c_int)]),
    "mvs_maps_release": (c_int, [c_void_p, c_int]),
    "mvs_presmooth": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_size_t, c_double, c_void_p, c_size_t]),
    "mvs_sgbm_check": (c_int, [C.POINTER(SgbmParams), c_int, c_int]),
    "mvs_set_sgbm_recurrence": (c_int, [c_void_p, c_int]),
    "mvs_sgbm_compute16": (c_int, [c_void_p, C.POINTER(SgbmParams), c_void_p, c_void_p, c_int, c_int, c_size_t, c_void_p,
                                   c_size_t]),
    "mvs_disparity_sgbm": (c_int, [c_void_p, C.POINTER(SgbmParams), c_double, c_void_p, c_void_p, c_int, c_int, c_int,
                                   c_size_t, c_void_p, c_size_t]),
    "mvs_disparity_census": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_size_t, c_int, c_void_p, c_size_t, c_void_p]),
    "mvs_disp_merge_ref": (c_int, [c_void_p, c_void_p, c_size_t, c_void_p, c_size_t, c_void_p, c_size_t, c_void_p, c_int, c_int, c_int]),
    "mvs_point_cloud": (c_int, [c_void_p, C.POINTER(Bcam), c_void_p, c_size_t, c_int, c_int, c_int, c_void_p, c_void_p,
                                c_void_p, C.POINTER(C.c_int32), c_void_p]),
    "mvs_point_cloud_static_dev": (c_int, [c_void_p, C.POINTER(Bcam), c_void_p, c_int, c_int, c_int, c_void_p, c_void_p,
                                           c_void_p, c_void_p, c_void_p]),
    "mvs_hist_dims": (c_int,[c_void_p, c_float, C.POINTER(c_int), C.POINTER(c_int)]),
    "mvs_floor_hist_camera": (c_int, [c_void_p, C.POINTER(HistParams), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                      C.c_int32, c_void_p, c_int, c_int, c_void_p, c_int, c_void_p, c_int, c_int]),
    "mvs_floor_hist_camera_cells": (c_int, [c_void_p, C.POINTER(HistParams), c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                            C.c_int32, c_void_p, c_int, c_int, c_void_p, c_int, c_void_p, c_int, c_int, c_void_p]),
    "mvs_hist_sum": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "mvs_hist_post": (c_int, [c_void_p, c_void_p, c_int, c_int, c_float, c_float, c_float]),
    "mvs_hist_post_dev": (c_int, [c_void_p, c_void_p, c_int, c_int, c_float, c_float, c_float]),
    "mvs_input_images": (c_int, [c_void_p, c_void_p, c_int, c_int, c_size_t, c_int, c_void_p, c_size_t, c_void_p, c_size_t]),
    "mvs_input_images_dev": (c_int, [c_void_p, c_void_p, c_int, c_int, c_size_t, c_int, c_void_p, c_size_t, c_void_p, c_size_t]),
    "mvs_localization_hists": (c_int, [c_void_p, C.POINTER(LocParams), c_void_p, c_int, c_int, c_double] + [c_void_p] * 5),
    "mvs_localization_hists_dev": (c_int, [c_void_p, C.POINTER(LocParams), c_void_p, c_int, c_int, c_double] + [c_void_p] * 5),
    "mvs_frame_run": (c_int, [c_void_p, c_int, C.POINTER(FrameDesc), C.POINTER(FrameResult)]),
    "mvs_frame_run_batch": (c_int, [c_void_p, c_int, C.POINTER(FrameDesc), c_int, C.POINTER(FrameResult)]),
    "mvs_frame_fetch": (c_int, [c_void_p, c_int, c_void_p, c_size_t, c_void_p, c_void_p, c_void_p, C.POINTER(C.c_int32),
                                c_void_p, c_void_p]),
    "mvs_frame_fetch_batch": (c_int, [c_void_p, c_int, c_int, C.POINTER(FetchDesc)]),
    "mvs_frame_hist_sum": (c_int, [c_void_p, c_int, c_int, C.POINTER(c_void_p), C.POINTER(c_int), C.POINTER(c_int)]),
    "mvs_sgbm_volumes": (c_int, [c_void_p, C.POINTER(SgbmParams), c_void_p, c_void_p, c_int, c_int, c_size_t, c_void_p,
                                 c_void_p, c_void_p]),
    "mvs_test_raise_wave_error": (c_int, [c_void_p, c_int]),
    "mvs_stage_ms": (c_float, [c_void_p, C.c_char_p]),
}

_lib = None


class MvsError(RuntimeError):
    pass


def load():
    """dlopen libmvstereo.so and set the prototypes.  Raises if the library is missing (no fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MvsError(f"{LIB_PATH} is missing: build it with `python -m multiview_b200.build` (there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
