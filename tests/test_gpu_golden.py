"""The CUDA path, through the C ABI, against the committed cv2 golden vectors (tests/golden/cv2_golden.npz)."""
import os

import numpy as np
import pytest

from multiview_b200 import api

pytestmark = pytest.mark.gpu

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "cv2_golden.npz"))
SGBM_NAMES = sorted(k[len("sgbm_"):-len("_disp16")] for k in G.files if k.startswith("sgbm_") and k.endswith("_disp16"))


def test_remap_golden(ctx):
    assert np.array_equal(api.calc_rectified(G["remap_src"], G["remap_mapx"], G["remap_mapy"], ctx=ctx), G["remap_out"])


def test_presmooth_golden(ctx):
    assert np.array_equal(api.presmooth(G["presmooth_in"], 2.15, ctx=ctx), G["presmooth_out"])


@pytest.mark.parametrize("name", SGBM_NAMES)
def test_sgbm_golden(ctx, name):
    p = api.StereoSGBMParams(*[int(v) for v in G[f"sgbm_{name}_params"]][:2], *[int(v) for v in G[f"sgbm_{name}_params"]][2:])
    got = api.StereoSGBMDisparity.compute16(p, G[f"sgbm_{name}_L"], G[f"sgbm_{name}_R"], ctx=ctx)
    assert np.array_equal(got, G[f"sgbm_{name}_disp16"])


def test_chain_golden(ctx):
    rect = [api.calc_rectified(G[f"chain_sensor{i}"], G[f"chain_mapx{i}"], G[f"chain_mapy{i}"], ctx=ctx) for i in range(2)]
    p = api.DisparityParams(stereosgbm_params=api.StereoSGBMParams(sg_num_disparities=64, sg_mode=1))
    got = api.Disparity.calculate(p, True, rect[0], rect[1], ctx=ctx)
    assert np.array_equal(got, G["chain_disp"])


GW = np.load(os.path.join(os.path.dirname(__file__), "golden", "cv2_golden_widened.npz"))


@pytest.mark.parametrize("interp", [0, 1, 2, 4])
def test_remap_modes_golden(ctx, interp):
    got = api.calc_rectified(GW["remap_src"], GW["remap_mapx"], GW["remap_mapy"], interp, ctx=ctx)
    assert np.array_equal(got, GW[f"remap_out_{interp}"])


@pytest.mark.parametrize("eq", [0, 1])
def test_input_images_golden(ctx, eq):
    equal, grey = api.input_images(GW["input_raw"], bool(eq), ctx=ctx)
    assert np.array_equal(equal, GW[f"input_equal_{eq}"]) and np.array_equal(grey, GW[f"input_grey_{eq}"])
