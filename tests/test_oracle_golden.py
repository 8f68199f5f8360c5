"""The oracle (C restatement) against the committed cv2 golden vectors -- CPU only."""
import os

import numpy as np
import pytest

import oracle

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "cv2_golden.npz"))
SGBM_NAMES = sorted(k[len("sgbm_"):-len("_disp16")] for k in G.files if k.startswith("sgbm_") and k.endswith("_disp16"))


def test_remap_golden():
    got = oracle.remap_cubic_bgr8(G["remap_src"], G["remap_mapx"], G["remap_mapy"])
    assert np.array_equal(got, G["remap_out"])
    assert ((G["remap_out"] == 255).all(axis=2)).mean() > 0.02  # the border colour is exercised


def test_presmooth_golden():
    assert np.array_equal(oracle.presmooth(G["presmooth_in"], 2.15), G["presmooth_out"])


@pytest.mark.parametrize("name", SGBM_NAMES)
def test_sgbm_golden(name):
    p = oracle.SgbmParams(*[int(v) for v in G[f"sgbm_{name}_params"]])
    got = oracle.sgbm_compute(p, G[f"sgbm_{name}_L"], G[f"sgbm_{name}_R"])
    assert np.array_equal(got, G[f"sgbm_{name}_disp16"])


def test_chain_golden():
    rect = [oracle.remap_cubic_bgr8(G[f"chain_sensor{i}"], G[f"chain_mapx{i}"], G[f"chain_mapy{i}"]) for i in range(2)]
    p = oracle.SgbmParams.reference_defaults(64, 1)
    got = oracle.disp_to_float(oracle.sgbm_compute(p, oracle.presmooth(rect[0]), oracle.presmooth(rect[1])))
    assert np.array_equal(got, G["chain_disp"])
    assert (got > 0).mean() > 0.3


def test_taps_are_consistent():
    """S1 (after the forward pass) + the median/speckle helpers reproduce the full result."""
    name = "hh_default"
    p = oracle.SgbmParams(*[int(v) for v in G[f"sgbm_{name}_params"]])
    t = oracle.sgbm_compute(p, G[f"sgbm_{name}_L"], G[f"sgbm_{name}_R"], taps=True)
    assert (t["S"] >= t["S1"]).all() and (t["C"] >= 0).all()
    d = oracle.filter_speckles(oracle.median3_s16(t["raw"]), (p.min_disp - 1) * 16, p.speckle_win, 16 * p.speckle_range)
    assert np.array_equal(d, t["disp"])


GW = np.load(os.path.join(os.path.dirname(__file__), "golden", "cv2_golden_widened.npz"))


@pytest.mark.parametrize("interp", [0, 1, 2, 4])
def test_remap_modes_golden(interp):
    got = oracle.remap_bgr8(GW["remap_src"], GW["remap_mapx"], GW["remap_mapy"], interp)
    assert np.array_equal(got, GW[f"remap_out_{interp}"])


@pytest.mark.parametrize("eq", [0, 1])
def test_input_images_golden(eq):
    equal, grey = oracle.input_images(GW["input_raw"], bool(eq))
    assert np.array_equal(equal, GW[f"input_equal_{eq}"]) and np.array_equal(grey, GW[f"input_grey_{eq}"])


def test_hist_post_golden():
    got = oracle.hist_post(GW["hist_in"], 0.5, 0.0, 1000.0)
    assert np.allclose(got, GW["hist_post"], rtol=1e-6, atol=1e-4)  # float blur: OpenCV's result varies by an ulp with its SIMD path
