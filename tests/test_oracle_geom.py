"""Sanity of the fp64 geometry restatement against independent numpy formulas -- CPU only.
(The reference has no runnable build here and no golden vectors for these functions: parity unpinned.)"""
import numpy as np

import oracle
from multiview_b200 import api, synth


def _bcam(w=320, h=200):
    K = synth.intrinsics(w, h)
    return oracle.bcam_init(synth.bcam_quaternion(), synth.BCAM_T, synth.BCAM_BASELINE, K), K


def test_solve3d_reprojects():
    b, K = _bcam()
    bc = api.BinocularCamera.from_c(b)
    KRR = [np.linalg.inv(m) for m in bc.KRR_inv]
    for (x, d, y) in [(100.0, 20.0, 50.0), (300.0, 3.5, 199.0), (10.0, 60.0, 0.0)]:
        X = oracle.solve3d(b, x, x - d, y)
        # larger disparity = closer point; depth along the rectified axis ~ f*B/d
        assert 0.5 < (KRR[0] @ X)[2] / (K[0, 0] * synth.BCAM_BASELINE / d) / np.linalg.norm(np.linalg.inv(K) @ KRR[0], 2) < 2.0
        for i, xi in enumerate((x, x - d)):
            p = KRR[i] @ (X - bc.C[i])
            assert np.allclose(p[:2] / p[2], (xi, y), atol=1e-6)


def test_point_cloud_order_and_lookup():
    b, _ = _bcam(64, 40)
    rng = np.random.default_rng(0)
    disp = rng.uniform(1, 30, (40, 64)).astype(np.float32)
    disp[rng.random(disp.shape) < 0.4] = 0
    disp[3, 3] = np.nan
    pc = oracle.point_cloud(b, disp)
    valid = np.isfinite(disp) & (disp != 0)
    assert pc["N"] == valid.sum()
    ys, xs = np.nonzero(valid)
    assert np.array_equal(pc["xy"], np.stack([xs, ys], 1))
    assert np.array_equal(pc["lookup"][valid], np.arange(pc["N"])) and (pc["lookup"][~valid] == -1).all()
    assert np.allclose(pc["C"], pc["Xs"].mean(0), rtol=1e-12)


def test_et_apply_is_rotation_plus_translation():
    import ctypes

    et = synth.camera_extrinsics(2, 5)
    q = np.array(et[3:7])
    x, y, z, w = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    X = np.array([0.3, -1.2, 4.0])
    Y = (ctypes.c_double * 3)()
    oracle.lib().orc_et_apply((ctypes.c_double * 8)(*et), (ctypes.c_double * 3)(*X), Y)
    assert np.allclose(np.array(Y[:]), R @ X + np.array(et[:3]), atol=1e-12)
    # the camera looks down at the floor: a point 4 m ahead lands below the camera
    assert Y[2] < et[2]


def test_floor_hist_counts_by_hand():
    hp = oracle.HistParams(0.5, 0.5, 2.0, 0, 0)
    et = (0, 0, 0, 0, 0, 0, 1, 1.0)
    aabb = (-1.0, -1.0, 1.0, 1.0)
    Xs = np.array([[0.0, 0.0, 1.0], [0.24, 0.0, 1.0], [0.26, 0.0, 1.0], [0.0, 0.0, 0.4], [0.0, 0.0, 2.0], [5.0, 0.0, 1.0],
                   [-1.0, -1.0, 0.5], [0.99, 0.99, 1.0]])
    xy = np.zeros((len(Xs), 2), np.int32)
    h = oracle.floor_hist_camera(hp, et, aabb, (0, 0, 0), Xs, xy)
    assert h.shape == (4, 4)
    want = np.zeros((4, 4), np.float32)
    want[2, 2] = 3  # (0,0), (0.24,0) round to bin 2; z = 2.0 is inside the closed interval
    want[2, 3] = 1  # 0.26 / 0.5 + 2 = 2.52 -> 3
    want[0, 0] = 1
    assert np.array_equal(h, want)  # (0.99, 0.99) rounds to bin 4 = out of bounds; z = 0.4 rejected; x = 5 rejected
    hp2 = oracle.HistParams(0.5, 0.5, 2.0, 1, 0)
    h2 = oracle.floor_hist_camera(hp2, et, aabb, (0, 0, 0), Xs[:1], xy[:1])
    assert h2[2, 2] == np.float32(1.0)
    assert np.array_equal(oracle.hist_sum(np.stack([h, h2])), h + h2)


def test_localization_front_half_against_numpy_formulas():
    """The C restatement of localization-data.cpp:147-315 against an independent numpy/scipy formulation."""
    from scipy.special import erfc

    rng = np.random.default_rng(5)
    hist = ((rng.random((40, 30)) < 0.4) * rng.gamma(2.0, 100.0, (40, 30))).astype(np.float32)
    p = oracle.LocParams.reference_defaults()
    hist_sz = 0.1
    out = oracle.localization_hists(p, hist, hist_sz)
    motion = np.where(hist.astype(np.float64) > p.n_deviations, hist, 0).astype(np.float32)
    assert np.array_equal(out["motion"], motion)
    radius = int(np.ceil(0.5 * p.person_diameter / hist_sz))
    rinv = float(np.float32(1.0) / (np.float32(0.5) * np.float32(0.5 * p.person_diameter)))
    want = np.zeros_like(motion, dtype=np.float64)
    for y0 in range(40):
        for x0 in range(30):
            vals = []
            for dy in range(-radius, radius + 1):
                for dx in range(-radius, radius + 1):
                    if dx * dx + dy * dy <= radius * radius and 0 <= x0 + dx < 30 and 0 <= y0 + dy < 40:
                        wg = 1.0 - np.sqrt(dx * dx + dy * dy) * hist_sz * rinv
                        if wg >= 0:
                            vals.append(float(np.float32(wg * wg * float(motion[y0 + dy, x0 + dx]))))
            want[y0, x0] = sum(vals) / len(vals) if vals else 0.0
    assert np.allclose(out["rhist"], want.astype(np.float32), rtol=1e-6)
    phi = lambda z: 0.5 * erfc(-z / np.sqrt(2.0))
    r = out["rhist"].astype(np.float64)
    assert np.allclose(out["bg"], np.minimum(p.prob_bg_max, phi(-(r - p.background_median) / p.background_stddev)), rtol=1e-6)
    assert np.allclose(out["fg0"], phi((r - p.track_median) / p.track_stddev), rtol=1e-6)
    assert np.allclose(out["fg"], np.maximum(out["fg0"], 0.1) ** 2, rtol=1e-6)
