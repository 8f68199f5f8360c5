"""Generates the golden fixtures in this directory from the OpenCV the reference calls (cv2 4.13.0 here).

    python tests/golden/make_golden.py

Run in the build container (needs cv2).  Inputs AND cv2 outputs are stored, so the tests never depend on
cv2 or on the random generator reproducing the inputs.  cv2 is driven single-threaded with IPP off, like the
reference's WITH_IPP=OFF build (build/install-files/build-scripts/build-opencv.sh:57).
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import oracle  # noqa: E402
from multiview_b200 import synth  # noqa: E402
from oracle import cv2ref  # noqa: E402

SGBM_CASES = [
    # name, W, H, D, (block, d12, minD, sw, sr, cap, ur, mode, p1, p2), noise
    ("hh_default", 200, 64, 64, (13, 596, 1, 600, 250, 54, 3, 1, 22, 1000), False),
    ("hh4_default", 200, 64, 64, (13, 596, 1, 600, 250, 54, 3, 3, 22, 1000), False),
    ("sgbm_default", 200, 64, 64, (13, 596, 1, 600, 250, 54, 3, 0, 22, 1000), False),
    ("hh_small_block", 160, 56, 32, (5, 2, 1, 0, 0, 54, 10, 1, 8, 32), False),
    ("hh_noise_saturated", 160, 56, 32, (13, 596, 1, 50, 2, 54, 3, 1, 22, 1000), True),
    ("sgbm_mind3", 160, 56, 32, (5, 1, 3, 20, 1, 31, 5, 0, 10, 120), False),
    ("hh4_d96", 240, 50, 96, (9, 596, 1, 0, 0, 54, 3, 3, 22, 1000), False),
    ("hh_d256", 420, 36, 256, (13, 596, 1, 600, 250, 54, 3, 1, 22, 1000), False),
]


def main():
    cv2ref.set_threads(1)
    rng = np.random.default_rng(20261106)
    out = {}
    # a1 remap
    sw, sh, w, h = 200, 150, 160, 120
    src = rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8)
    mx, my = synth.radial_maps(w, h, 0.3, (1.5, -2.25))
    mx = (mx * sw / w + rng.normal(0, 0.3, (h, w))).astype(np.float32)
    my = (my * sh / h + rng.normal(0, 0.3, (h, w))).astype(np.float32)
    out["remap_src"], out["remap_mapx"], out["remap_mapy"] = src, mx, my
    out["remap_out"] = cv2ref.remap_cubic_bgr8(src, mx, my)
    # a2 grey + double blur
    img = rng.integers(0, 256, (97, 131, 3), dtype=np.uint8)
    out["presmooth_in"] = img
    out["presmooth_out"] = cv2ref.presmooth(img, 2.15)
    # a3 StereoSGBM
    for name, W, H, D, prm, noise in SGBM_CASES:
        if noise:
            L = rng.integers(0, 256, (H, W), dtype=np.uint8)
            R = rng.integers(0, 256, (H, W), dtype=np.uint8)
        else:
            l, r = synth.rectified_pair(W, H, D, 77 + D)
            L, R = cv2ref.presmooth(l, 0.0), cv2ref.presmooth(r, 0.0)
        blk, d12, minD, sw_, sr, cap, ur, mode, p1, p2 = prm
        p = oracle.SgbmParams(blk, d12, minD, D, sw_, sr, cap, ur, mode, p1, p2)
        out[f"sgbm_{name}_L"], out[f"sgbm_{name}_R"] = L, R
        out[f"sgbm_{name}_params"] = np.array([blk, d12, minD, D, sw_, sr, cap, ur, mode, p1, p2], np.int32)
        out[f"sgbm_{name}_disp16"] = cv2ref.sgbm_compute(p, L, R)
    # a1-a4 chain: sensor images -> float disparity (Disparity::calculate)
    cam = synth.make_camera(320, 200, 64, config=9)
    rect = [cv2ref.remap_cubic_bgr8(cam["sensor"][i], cam["mapx"][i], cam["mapy"][i]) for i in range(2)]
    p = oracle.SgbmParams.reference_defaults(64, 1)
    for i in range(2):
        out[f"chain_sensor{i}"], out[f"chain_mapx{i}"], out[f"chain_mapy{i}"] = cam["sensor"][i], cam["mapx"][i], cam["mapy"][i]
    out["chain_disp"] = cv2ref.disparity_calculate(p, rect[0], rect[1], 2.15)
    np.savez_compressed(os.path.join(HERE, "cv2_golden.npz"), **out)
    print("wrote", os.path.join(HERE, "cv2_golden.npz"), {k: v.shape for k, v in out.items() if "disp" in k or k.endswith("_out")})


def main_widened():
    """Second fixture file (rows added later in round 1): the other cv::remap modes, the input_images node and the
    float histogram blur.  Own seed, so the first file never changes."""
    cv2ref.set_threads(1)
    rng = np.random.default_rng(20261107)
    out = {}
    sw, sh, w, h = 120, 90, 96, 72
    src = rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8)
    mx = (rng.random((h, w)) * (sw + 12) - 6).astype(np.float32)
    my = (rng.random((h, w)) * (sh + 12) - 6).astype(np.float32)
    mx[0, 0], my[0, 1], mx[1, 1], my[2, 2] = 1e9, -1e9, np.nan, 3e4
    out["remap_src"], out["remap_mapx"], out["remap_mapy"] = src, mx, my
    for interp in (0, 1, 2, 4):
        out[f"remap_out_{interp}"] = cv2ref.remap_bgr8(src, mx, my, interp)
    img = synth.make_camera(320, 200, 64, config=3)["sensor"][0]
    out["input_raw"] = img
    for eq in (0, 1):
        e, g = cv2ref.input_images(img, bool(eq))
        out[f"input_equal_{eq}"], out[f"input_grey_{eq}"] = e, g
    hist = ((rng.random((100, 120)) < 0.2) * rng.integers(1, 1500, (100, 120))).astype(np.float32)
    out["hist_in"] = hist
    out["hist_post"] = cv2ref.hist_post(hist, 0.5, 0.0, 1000.0)
    np.savez_compressed(os.path.join(HERE, "cv2_golden_widened.npz"), **out)
    print("wrote", os.path.join(HERE, "cv2_golden_widened.npz"), sorted(out))


if __name__ == "__main__":
    if "--widened-only" not in sys.argv:
        main()
    main_widened()
