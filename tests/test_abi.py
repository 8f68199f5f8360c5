"""The C-ABI library loads and exports every symbol include/mvstereo.h declares -- CPU only, no compute."""
import ctypes
import os
import re

import pytest

from multiview_b200 import _lib, api, build

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build()
    return _lib.load()


def _declared():
    txt = open(os.path.join(ROOT, "include", "mvstereo.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(mvs_[a-z0-9_]+)\s*\(", txt)))


def test_every_declared_symbol_is_exported(lib):
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in mvstereo.h but not exported"
    assert set(names) == set(_lib.SIGNATURES), set(names) ^ set(_lib.SIGNATURES)


def test_version_and_param_check(lib):
    assert b"sm_100a" in lib.mvs_version()
    ok = api.StereoSGBMParams().to_c()
    assert lib.mvs_sgbm_check(ctypes.byref(ok), 896, 672) == 0
    for kw in (dict(sg_SAD_window_size=21), dict(sg_p2=30000), dict(sg_num_disparities=100), dict(sg_min_disparity=-4),
               dict(sg_mode=2), dict(sg_num_disparities=512)):
        bad = api.StereoSGBMParams(**kw).to_c()
        assert lib.mvs_sgbm_check(ctypes.byref(bad), 896, 672) < 0, kw
    assert lib.mvs_sgbm_check(ctypes.byref(ok), 90, 60) < 0  # narrower than the disparity range


def test_hist_dims_matches_reference_formula(lib):
    # ensure_floor_hist_dims, floor-histogram.cpp:19-35; hist_sz is a float promoted to double
    assert api.hist_dims((-5.0, -5.0, 5.0, 5.0), 0.10) == (100, 100)
    assert api.hist_dims((-8.0, -8.0, 8.0, 8.0), 0.10) == (160, 160)
    assert api.hist_dims((0.0, 0.0, 1.05, 2.0), 0.25) == (5, 8)
    import oracle

    for aabb, sz in [((-5.0, -5.0, 5.0, 5.0), 0.1), ((-3.3, -1.2, 7.9, 4.4), 0.07), ((0, 0, 1, 1), 0.3)]:
        assert api.hist_dims(aabb, sz) == oracle.hist_dims(aabb, sz)


def test_no_cpu_fallback(lib):
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = ctypes.c_void_p()
    assert lib.mvs_ctx_create(0, ctypes.byref(h)) < 0 and not h.value
    with pytest.raises(api.MvsError):
        api.Context(0)
