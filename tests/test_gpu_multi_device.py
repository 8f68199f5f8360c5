"""Two contexts on two devices inside ONE process (the C ABI's per-device one-time state); skipped with fewer than 2 GPUs."""
import numpy as np
import pytest

import oracle
from multiview_b200 import api, synth

pytestmark = pytest.mark.gpu


def test_contexts_on_two_devices_give_the_same_frame():
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    W, H, D = 320, 200, 64
    cam = synth.make_camera(W, H, D, config=5)
    sp = api.StereoSGBMParams(sg_num_disparities=D, sg_mode=1)
    hp = api.FloorHistogramParams(weighting=api.WGT_NONE, apply_slic_still_weights=False)
    res = []
    for dev in (1, 0, 1):  # the second device first: its kernels need their attributes set there too
        ctx = api.Context(dev)
        desc = api.make_frame_desc(cam, sp, hp, synth.HIST_AABB)
        api.frame_run(ctx, 0, desc)
        out = api.frame_fetch(ctx, 0, W, H, *api.hist_dims(synth.HIST_AABB, hp.hist_sz))
        res.append(out)
    for k in ("disp", "cam_hist", "lookup"):
        assert np.array_equal(res[0][k], res[1][k]) and np.array_equal(res[0][k], res[2][k])
    rect = [oracle.remap_cubic_bgr8(cam["sensor"][i], cam["mapx"][i], cam["mapy"][i]) for i in range(2)]
    want = oracle.disp_to_float(oracle.sgbm_compute(oracle.SgbmParams.reference_defaults(D, 1), oracle.presmooth(rect[0]), oracle.presmooth(rect[1])))
    assert np.array_equal(res[0]["disp"], want)
