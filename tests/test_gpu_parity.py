"""GPU parity tests: every stage of the CUDA path, called through the C ABI, against the CPU oracle.

Bit-exact for integer/byte/index results; point coordinates within 1e-4 relative (north_star)."""
import ctypes
import ctypes as C

import numpy as np
import pytest

import oracle
from multiview_b200 import api, synth

pytestmark = pytest.mark.gpu


def _pair(W, H, D, seed, noise=False):
    if noise:
        rng = np.random.default_rng(seed)
        return rng.integers(0, 256, (H, W), dtype=np.uint8), rng.integers(0, 256, (H, W), dtype=np.uint8)
    l, r = synth.rectified_pair(W, H, D, seed)
    return oracle.bgr2gray(l), oracle.bgr2gray(r)


def _sp(D, mode, block=13, p1=22, p2=1000, ur=3, d12=596, sw=600, sr=250, cap=54, minD=1):
    return api.StereoSGBMParams(block, d12, minD, D, sw, sr, cap, ur, mode, p1, p2)


def _op(sp):
    return oracle.SgbmParams(sp.sg_SAD_window_size, sp.sg_12_max_diff, sp.sg_min_disparity,
                                                     sp.sg_num_disparities, sp.sg_speckle_window_size, sp.sg_speckle_range,
                                                     sp.sg_prefilter_cap, sp.sg_uniqueness_ratio, sp.sg_mode, sp.sg_p1, sp.sg_p2)


def test_remap_cubic_bit_exact(ctx):
    rng = np.random.default_rng(7)
    for (sw, sh, w, h) in [(128, 96, 128, 96), (200, 150, 160, 120), (64, 48, 80, 60)]:
        src = rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8)
        mx, my = synth.radial_maps(w, h, 0.3, (1.5, -2.25))
        mx = (mx * sw / w + rng.normal(0, 0.3, (h, w))).astype(np.float32)
        my = (my * sh / h + rng.normal(0, 0.3, (h, w))).astype(np.float32)
        mx[0, 0] = 1e9
        my[0, 1] = -1e9
        mx[1, 1] = np.nan
        got = api.calc_rectified(src, mx, my, ctx=ctx)
        want = oracle.remap_cubic_bgr8(src, mx, my)
        assert np.array_equal(got, want), f"remap mismatch {(got != want).sum()} px"


@pytest.mark.parametrize("interp", [api.INTER_NEAREST, api.INTER_LINEAR, api.INTER_CUBIC, api.INTER_LANCZOS4])
def test_remap_interpolation_modes_bit_exact(ctx, interp):
    """calc_rectified.hpp:17-22: every interpolation mode of the node, against the cv2-pinned oracle."""
    rng = np.random.default_rng(70 + interp)
    for (sw, sh, w, h) in [(80, 60, 70, 50), (33, 21, 64, 40), (200, 150, 160, 120)]:
        src = rng.integers(0, 256, (sh, sw, 3), dtype=np.uint8)
        mx = (rng.random((h, w)) * (sw + 12) - 6).astype(np.float32)
        my = (rng.random((h, w)) * (sh + 12) - 6).astype(np.float32)
        mx[0, 0], my[0, 1], mx[1, 1], my[2, 2] = 1e9, -1e9, np.nan, 3e4
        mx[3, :8] = np.arange(8, dtype=np.float32) * 0.5
        got = api.calc_rectified(src, mx, my, interp, ctx=ctx)
        want = oracle.remap_bgr8(src, mx, my, interp)
        assert np.array_equal(got, want), f"interp {interp}: {(got != want).any(axis=2).sum()} px differ"
    with pytest.raises(api.MvsError):
        api.calc_rectified(src, mx, my, 3, ctx=ctx)  # INTER_AREA is not a remap mode


def test_presmooth_bit_exact(ctx):
    rng = np.random.default_rng(8)
    # widths that are / are not multiples of 4 (packed and scalar kernels), every tap-radius class of the packed kernels
    for (w, h) in [(131, 97), (20, 12), (640, 48), (64, 40), (36, 36)]:
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        for sigma in (2.15, 1.0, 5.0, 3.2, 0.4, 0.8):
            assert np.array_equal(api.presmooth(img, sigma, ctx=ctx), oracle.presmooth(img, sigma))
        assert np.array_equal(api.presmooth(img[..., 0].copy(), 2.15, ctx=ctx), oracle.presmooth(img[..., 0].copy(), 2.15))
        assert np.array_equal(api.presmooth(img, 0.0, ctx=ctx), oracle.bgr2gray(img))


@pytest.mark.parametrize("mode", [1, 0, 3])
@pytest.mark.parametrize("W,H,D", [(72, 40, 16), (200, 64, 64), (240, 50, 96), (300, 48, 128), (420, 36, 256)])
def test_sgbm_volumes_and_disparity(ctx, mode, W, H, D):
    for noise in (False, True):
        L, R = _pair(W, H, D, 100 + D + mode, noise)
        sp = _sp(D, mode)
        t = oracle.sgbm_compute(_op(sp), L, R, taps=True)
        Cg, Sg, rawg = api.StereoSGBMDisparity.volumes(sp, L, R, ctx=ctx)
        assert np.array_equal(Cg, t["C"]), f"C mismatch {(Cg != t['C']).sum()}"
        assert np.array_equal(Sg, t["S"]), f"S mismatch {(Sg != t['S']).sum()}"
        assert np.array_equal(rawg, t["raw"]), f"raw disparity mismatch {(rawg != t['raw']).sum()}"
        d16 = api.StereoSGBMDisparity.compute16(sp, L, R, ctx=ctx)
        assert np.array_equal(d16, t["disp"]), f"disp16 mismatch {(d16 != t['disp']).sum()}"


@pytest.mark.parametrize("params", [
    dict(block=1, p1=4, p2=16, ur=0, d12=1, sw=0, sr=0),
    dict(block=3, p1=8, p2=32, ur=0, d12=1, sw=0, sr=0),
    dict(block=5, p1=8, p2=32, ur=10, d12=2, sw=0, sr=0),
    dict(block=9, p1=22, p2=1000, ur=3, d12=596, sw=0, sr=0),
    dict(block=13, p1=22, p2=1000, ur=25, d12=1, sw=50, sr=2),
    dict(block=11, p1=22, p2=4000, ur=3, d12=596, sw=0, sr=0),
    dict(block=7, p1=0, p2=0, ur=-1, d12=0, sw=100, sr=1, cap=10),
    dict(block=5, p1=10, p2=120, ur=5, d12=1, sw=0, sr=0, minD=0),
    dict(block=5, p1=10, p2=120, ur=5, d12=1, sw=20, sr=1, minD=3),
])
def test_sgbm_parameter_matrix(ctx, params):
    for mode in (1, 0, 3):
        for noise in (False, True):
            L, R = _pair(160, 56, 32, 5 + mode, noise)
            sp = _sp(32, mode, **params)
            want = oracle.sgbm_compute(_op(sp), L, R)
            got = api.StereoSGBMDisparity.compute16(sp, L, R, ctx=ctx)
            assert np.array_equal(got, want), f"mode {mode} noise {noise}: {(got != want).sum()} px differ"


def test_disparity_calculate_float(ctx):
    cam = synth.make_camera(320, 200, 64, config=9)
    rect = [oracle.remap_cubic_bgr8(cam["sensor"][i], cam["mapx"][i], cam["mapy"][i]) for i in range(2)]
    p = api.DisparityParams(stereosgbm_params=_sp(64, 1))
    got = api.Disparity.calculate(p, True, rect[0], rect[1], ctx=ctx)
    want = oracle.disp_to_float(oracle.sgbm_compute(_op(p.stereosgbm_params), oracle.presmooth(rect[0]), oracle.presmooth(rect[1])))
    assert got.dtype == np.float32 and np.array_equal(got, want)
    assert (got > 0).mean() > 0.3


def test_unsupported_parameters_fail_loudly(ctx):
    L, R = _pair(160, 56, 32, 1)
    for bad in (dict(block=21), dict(p2=30000), dict(block=4)):
        with pytest.raises(api.MvsError):
            api.StereoSGBMDisparity.compute16(_sp(32, 1, **bad), L, R, ctx=ctx)
    with pytest.raises(api.MvsError):
        api.StereoSGBMDisparity.compute16(_sp(48, 1), L[:, :40].copy(), R[:, :40].copy(), ctx=ctx)  # too narrow
    with pytest.raises(api.MvsError):
        api.StereoSGBMDisparity.compute16(_sp(32, 2), L, R, ctx=ctx)  # MODE_SGBM_3WAY not on the path


def _cloud_inputs(w=320, h=200, D=64, seed=3):
    rng = np.random.default_rng(seed)
    disp = rng.uniform(2.0, 0.8 * D, (h, w)).astype(np.float32)
    disp[rng.random((h, w)) < 0.3] = 0.0
    disp[0, :5] = [np.nan, np.inf, -np.inf, -3.5, 0.0]
    K = synth.intrinsics(w, h)
    return disp, synth.bcam_quaternion(), synth.BCAM_T, synth.BCAM_BASELINE, K


def test_point_cloud_parity(ctx):
    disp, q, t, bl, K = _cloud_inputs()
    ob = oracle.bcam_init(q, t, bl, K)
    want = oracle.point_cloud(ob, disp)
    bc = api.BinocularCamera.from_c(ob)  # identical 24 doubles on both sides
    pc = api.PointCloud().init(disp, True, bc, ctx=ctx)
    assert pc.N() == want["N"]
    assert np.array_equal(pc.xy, want["xy"]) and np.array_equal(pc.lookup, want["lookup"])
    rel = np.abs(pc.Xs - want["Xs"]).max() / np.abs(want["Xs"]).max()
    assert rel <= 1e-4, rel  # north_star tolerance; the no-FMA build is expected to give 0
    assert np.allclose(pc.C, want["C"], rtol=1e-4, atol=0)
    # host-side BinocularCamera.init (numpy) agrees with the oracle's restatement
    bc2 = api.BinocularCamera().init(q, t, bl, K)
    assert np.allclose(bc2.KRR_inv[1], bc.KRR_inv[1], rtol=1e-9, atol=1e-12)
    # empty cloud: N == 0, centroid NaN like the reference's 0/0
    z = api.PointCloud().init(np.zeros((8, 8), np.float32), True, bc, ctx=ctx)
    assert z.N() == 0 and np.isnan(z.C).all() and (z.lookup == -1).all()


def test_floor_histogram_parity(ctx):
    disp, q, t, bl, K = _cloud_inputs()
    ob = oracle.bcam_init(q, t, bl, K)
    want_pc = oracle.point_cloud(ob, disp)
    aabb = synth.HIST_AABB
    n_cams = 3
    ets = [synth.camera_extrinsics(c, n_cams) for c in range(n_cams)]
    pcs = []
    for c in range(n_cams):
        pc = api.PointCloud()
        pc.Xs, pc.xy = want_pc["Xs"] * (1.0 + 0.1 * c), want_pc["xy"]
        pcs.append(pc)
    rng = np.random.default_rng(4)
    labels = rng.integers(-1, 50, disp.shape).astype(np.int32)
    still = rng.random(40)
    for weighting, use_still in [(api.WGT_NONE, False), (api.WGT_NORM, False), (api.WGT_QUADRANCE, True), (api.WGT_NORM, True)]:
        p = api.FloorHistogramParams(weighting=weighting, apply_slic_still_weights=use_still)
        fh = api.FloorHistogram()
        api.FloorHistogram.calculate(p, aabb, n_cams, ets, pcs, [e[:3] for e in ets], fh,
                                     slic_lookups=[labels] * n_cams if use_still else None,
                                     still_scores=[still] * n_cams if use_still else None, ctx=ctx)
        hp = oracle.HistParams(p.hist_sz, p.hist_min_z, p.hist_max_z, weighting, int(use_still))
        want = [oracle.floor_hist_camera(hp, ets[c], aabb, ets[c][:3], pcs[c].Xs, pcs[c].xy,
                                         labels if use_still else None, still if use_still else None) for c in range(n_cams)]
        for c in range(n_cams):
            assert (want[c] > 0).sum() > 10
            if weighting == api.WGT_NONE and not use_still:
                assert np.array_equal(fh.camera_hists[c], want[c])  # counts: bit-exact
            else:
                assert np.allclose(fh.camera_hists[c], want[c], rtol=1e-5, atol=1e-3)
        wsum = oracle.hist_sum(np.stack(want))
        if weighting == api.WGT_NONE and not use_still:
            assert np.array_equal(fh.hist_raw, wsum)
            # smoothing + clip (floor-histogram.cpp:411-434): same float operations in the same order -> identical
            assert np.array_equal(fh.hist, oracle.hist_post(wsum, p.gaussian_sigma, p.min_hist_count, p.max_hist_count))
        else:
            assert np.allclose(fh.hist_raw, wsum, rtol=1e-5, atol=1e-3)
            assert np.allclose(fh.hist, oracle.hist_post(wsum, p.gaussian_sigma, p.min_hist_count, p.max_hist_count), rtol=1e-5, atol=1e-2)


def test_floor_histogram_ordered_accumulation_is_bit_exact(ctx):
    """mvs_hist_params.ordered: every bin summed in point order like the reference's sequential loop, so the weighted
    histograms (the reference default is WGT_NORM with still weights) match the oracle bit for bit."""
    disp, q, t, bl, K = _cloud_inputs()
    want_pc = oracle.point_cloud(oracle.bcam_init(q, t, bl, K), disp)
    aabb = synth.HIST_AABB
    n_cams = 2
    ets = [synth.camera_extrinsics(c, n_cams) for c in range(n_cams)]
    pcs = []
    for c in range(n_cams):
        pc = api.PointCloud()
        pc.Xs, pc.xy = want_pc["Xs"] * (1.0 + 0.1 * c), want_pc["xy"]
        pcs.append(pc)
    rng = np.random.default_rng(14)
    labels = rng.integers(-1, 50, disp.shape).astype(np.int32)
    still = rng.random(40)
    for weighting, use_still in [(api.WGT_NORM, True), (api.WGT_QUADRANCE, False), (api.WGT_NONE, True)]:
        p = api.FloorHistogramParams(weighting=weighting, apply_slic_still_weights=use_still, ordered=True)
        fh = api.FloorHistogram()
        api.FloorHistogram.calculate(p, aabb, n_cams, ets, pcs, [e[:3] for e in ets], fh,
                                     slic_lookups=[labels] * n_cams if use_still else None,
                                     still_scores=[still] * n_cams if use_still else None, ctx=ctx)
        hp = oracle.HistParams(p.hist_sz, p.hist_min_z, p.hist_max_z, weighting, int(use_still))
        want = [oracle.floor_hist_camera(hp, ets[c], aabb, ets[c][:3], pcs[c].Xs, pcs[c].xy,
                                         labels if use_still else None, still if use_still else None) for c in range(n_cams)]
        for c in range(n_cams):
            assert (want[c] > 0).sum() > 10 and want[c].max() > 50
            assert np.array_equal(fh.camera_hists[c], want[c])
        assert np.array_equal(fh.hist_raw, oracle.hist_sum(np.stack(want)))


def test_frame_path_matches_stagewise_oracle(ctx):
    W, H, D = 384, 240, 64
    cam = synth.make_camera(W, H, D, config=2, cam=1, n_cams=4)
    sp = _sp(D, 1)
    hp = api.FloorHistogramParams(weighting=api.WGT_NONE, apply_slic_still_weights=False)
    desc = api.make_frame_desc(cam, sp, hp, synth.HIST_AABB)
    res = api.frame_run(ctx, 0, desc)
    out = api.frame_fetch(ctx, 0, W, H, res.hw, res.hh)
    rect = [oracle.remap_cubic_bgr8(cam["sensor"][i], cam["mapx"][i], cam["mapy"][i]) for i in range(2)]
    want_d = oracle.disp_to_float(oracle.sgbm_compute(_op(sp), oracle.presmooth(rect[0]), oracle.presmooth(rect[1])))
    assert np.array_equal(out["disp"], want_d)
    ob = oracle.bcam_init(cam["q"], cam["t"], cam["baseline"], cam["K"])
    # same calibration doubles on both sides: take them from the descriptor the GPU used
    ob2 = oracle.Bcam()
    for i in range(2):
        for k in range(9):
            ob2.krr_inv[i][k] = desc.bcam.krr_inv[i][k]
        for k in range(3):
            ob2.c[i][k] = desc.bcam.c[i][k]
    wpc = oracle.point_cloud(ob2, want_d)
    assert out["N"] == wpc["N"] and np.array_equal(out["xy"], wpc["xy"]) and np.array_equal(out["lookup"], wpc["lookup"])
    assert np.abs(out["Xs"] - wpc["Xs"]).max() <= 1e-4 * np.abs(wpc["Xs"]).max()
    ohp = oracle.HistParams(hp.hist_sz, hp.hist_min_z, hp.hist_max_z, 0, 0)
    wh = oracle.floor_hist_camera(ohp, cam["et"], synth.HIST_AABB, cam["et"][:3], wpc["Xs"], wpc["xy"])
    assert np.array_equal(out["cam_hist"], wh) and wh.sum() > 100
    # the reference's default weighting (distance to the camera) through the frame path, summed in point order
    hp2 = api.FloorHistogramParams(weighting=api.WGT_NORM, apply_slic_still_weights=False, ordered=True)
    res2 = api.frame_run(ctx, 1, api.make_frame_desc(cam, sp, hp2, synth.HIST_AABB))
    out2 = api.frame_fetch(ctx, 1, W, H, res2.hw, res2.hh)
    ohp2 = oracle.HistParams(hp2.hist_sz, hp2.hist_min_z, hp2.hist_max_z, api.WGT_NORM, 0)
    wh2 = oracle.floor_hist_camera(ohp2, cam["et"], synth.HIST_AABB, cam["et"][:3], out2["Xs"], out2["xy"])
    assert np.array_equal(out2["cam_hist"], wh2) and wh2.sum() > wh.sum()
    # oracle's own calibration restatement agrees with the host-side numpy one to rounding
    assert np.allclose(np.array(ob.krr_inv[1][:]), np.array(ob2.krr_inv[1][:]), rtol=1e-9, atol=1e-12)


def test_frame_batch_equals_single_frames(ctx):
    """mvs_frame_run_batch (one aggregation launch for all frames) gives exactly what per-frame calls give."""
    W, H, D = 352, 150, 64
    sp = _sp(D, 1)
    hp = api.FloorHistogramParams(weighting=api.WGT_NONE, apply_slic_still_weights=False)
    cams = [synth.make_camera(W, H, D, config=3, cam=c, n_cams=5) for c in range(5)]
    descs = [api.make_frame_desc(c, sp, hp, synth.HIST_AABB) for c in cams]
    single = []
    for i, d in enumerate(descs):
        res = api.frame_run(ctx, 10 + i, d)
        single.append(api.frame_fetch(ctx, 10 + i, W, H, res.hw, res.hh))
    res = api.frame_run_batch(ctx, 0, descs)
    for i in range(len(descs)):
        out = api.frame_fetch(ctx, i, W, H, res[i].hw, res[i].hh)
        assert np.array_equal(out["disp"], single[i]["disp"]) and (out["disp"] > 0).mean() > 0.3
        assert out["N"] == single[i]["N"] and np.array_equal(out["xy"], single[i]["xy"])
        assert np.array_equal(out["Xs"], single[i]["Xs"]) and np.array_equal(out["cam_hist"], single[i]["cam_hist"])
    # and the first one against the oracle
    rect = [oracle.remap_cubic_bgr8(cams[0]["sensor"][i], cams[0]["mapx"][i], cams[0]["mapy"][i]) for i in range(2)]
    want = oracle.disp_to_float(oracle.sgbm_compute(_op(sp), oracle.presmooth(rect[0]), oracle.presmooth(rect[1])))
    assert np.array_equal(single[0]["disp"], want)


def test_disp_map_update_merge(ctx):
    """a8: the reference-disparity merge of disp_map_update (holes and non-still super-pixels are filled)."""
    rng = np.random.default_rng(12)
    h, w = 90, 140
    disp = rng.uniform(-2, 60, (h, w)).astype(np.float32)
    disp[rng.random((h, w)) < 0.2] = 0.0
    disp[rng.random((h, w)) < 0.05] = np.nan
    ref = rng.uniform(1, 60, (h, w)).astype(np.float32)
    ref[rng.random((h, w)) < 0.1] = np.nan
    ref[rng.random((h, w)) < 0.02] = np.inf
    labels = rng.integers(-1, 40, (h, w)).astype(np.int32)  # some labels are beyond the 32-entry table
    cut = (rng.random(32) < 0.5).astype(np.uint8)
    got = api.disp_map_update(disp, ref, labels, cut, ctx=ctx)
    want = oracle.disp_merge_ref(disp, ref, labels, cut)
    assert np.array_equal(got, want, equal_nan=True)
    assert (got != disp).sum() > 1000 and np.array_equal(got[labels < 0], disp[labels < 0], equal_nan=True)


def test_sgbm_seeded_random_configurations(ctx):
    """Seeded sweep over sizes, disparity ranges, block sizes, penalties and modes inside the supported domain
    (mvs_sgbm_check): the whole SGBM chain against the cv2-pinned oracle, bit for bit."""
    rng = np.random.default_rng(2026)
    done = 0
    for _ in range(60):
        D = int(rng.choice([16, 32, 48, 64, 80, 96, 128, 160, 208, 256]))
        minD = int(rng.integers(0, 5))
        W = D + minD + int(rng.integers(8, 260))
        H = int(rng.integers(1, 72))
        block = int(rng.choice([1, 3, 5, 7, 9, 11, 13]))
        cap = int(rng.choice([15, 31, 54, 63]))
        p1 = int(rng.integers(0, 60))
        p2 = p1 + 1 + int(rng.integers(0, 3000))
        mode = int(rng.choice([0, 1, 3]))
        sp = _sp(D, mode, block=block, p1=p1, p2=p2, ur=int(rng.integers(0, 40)), d12=int(rng.choice([-1, 0, 1, 2, 596])),
                 sw=int(rng.choice([0, 30, 600])), sr=int(rng.choice([1, 2, 250])), cap=cap, minD=minD)
        if ctx.lib.mvs_sgbm_check(ctypes.byref(sp.to_c()), W, H) != 0:
            continue
        L, R = _pair(W, H, D, 9000 + done, noise=bool(rng.integers(0, 2)) or H < 16 or W < D + 48)
        want = oracle.sgbm_compute(_op(sp), L, R)
        got = api.StereoSGBMDisparity.compute16(sp, L, R, ctx=ctx)
        assert np.array_equal(got, want), f"W{W} H{H} D{D} minD{minD} block{block} mode{mode} p1 {p1} p2 {p2}: {(got != want).sum()} px differ"
        done += 1
    assert done >= 25


def test_localization_front_half(ctx):
    """motion / radius-filtered / bg / fg histograms from the floor histogram (localization-data.cpp:845-858)."""
    rng = np.random.default_rng(77)
    for (hh, hw, hist_sz, diam) in [(100, 100, 0.10, 0.4), (37, 53, 0.05, 0.4), (64, 48, 0.10, 0.9), (9, 7, 0.25, 0.4)]:
        hist = ((rng.random((hh, hw)) < 0.3) * rng.gamma(2.0, 120.0, (hh, hw))).astype(np.float32)
        p = api.LocalizationParams(person_diameter=diam)
        got = api.LocalizationHists.calculate(p, hist, hist_sz, ctx=ctx)
        op = oracle.LocParams(p.n_deviations, p.person_diameter, p.background_median, p.background_stddev, p.track_median,
                              p.track_stddev, p.prob_bg_max)
        want = oracle.localization_hists(op, hist, hist_sz)
        assert np.array_equal(got["motion"], want["motion"])
        assert np.array_equal(got["rhist"], want["rhist"])          # same fp64 operations in the same order
        for k in ("bg", "fg0", "fg"):                                  # erfc: CUDA vs libm, <= 1 ulp in double
            assert np.allclose(got[k], want[k], rtol=1e-6, atol=1e-12), k
            assert (got[k] != want[k]).mean() < 0.01, k


def test_input_images_node(ctx):
    """hist_equalize + BGR2GRAY (input_images.cpp:65-76), bit-exact against the cv2-pinned oracle."""
    rng = np.random.default_rng(31)
    imgs = [rng.integers(0, 256, (97, 131, 3), dtype=np.uint8), rng.integers(90, 140, (64, 80, 3), dtype=np.uint8),
            np.full((20, 24, 3), 77, np.uint8), synth.make_camera(640, 400, 64, config=3)["sensor"][0]]
    for img in imgs:
        for eq in (True, False):
            got, want = api.input_images(img, eq, ctx=ctx), oracle.input_images(img, eq)
            assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1])


def test_frame_batch_with_mixed_sensor_sizes_and_blur(ctx):
    """A batch whose frames differ in sensor size and presmooth sigma takes the per-image kernels instead of the
    one-launch-per-batch ones: results must still equal the per-frame calls."""
    W, H, D = 352, 150, 64
    sp = _sp(D, 1)
    hp = api.FloorHistogramParams(weighting=api.WGT_NONE, apply_slic_still_weights=False)
    cams = [synth.make_camera(W, H, D, config=4, cam=c, n_cams=3) for c in range(3)]
    # frame 1: sensor images padded to a wider, taller buffer (same maps: they address the original area)
    pad = [np.ascontiguousarray(np.pad(s, ((0, 6), (0, 10), (0, 0)), mode="edge")) for s in cams[1]["sensor"]]
    sigmas = [2.15, 2.15, 1.0]
    descs = []
    for i, c in enumerate(cams):
        d = api.make_frame_desc(c, sp, hp, synth.HIST_AABB, sigma=sigmas[i])
        if i == 1:
            for k in range(2):
                d.sensor[k] = pad[k].ctypes.data
            d.sw, d.sh, d.sstep = W + 10, H + 6, (W + 10) * 3
        descs.append(d)
    single = []
    for i, d in enumerate(descs):
        res = api.frame_run(ctx, 20 + i, d)
        single.append(api.frame_fetch(ctx, 20 + i, W, H, res.hw, res.hh))
    res = api.frame_run_batch(ctx, 0, descs)
    for i in range(len(descs)):
        out = api.frame_fetch(ctx, i, W, H, res[i].hw, res[i].hh)
        assert np.array_equal(out["disp"], single[i]["disp"]) and (out["disp"] > 0).mean() > 0.3
        assert out["N"] == single[i]["N"] and np.array_equal(out["cam_hist"], single[i]["cam_hist"])
    # the blur strength matters (frame 2 would differ with the default sigma)
    d2 = api.make_frame_desc(cams[2], sp, hp, synth.HIST_AABB)
    api.frame_run(ctx, 30, d2)
    assert not np.array_equal(api.frame_fetch(ctx, 30, W, H, res[0].hw, res[0].hh)["disp"], single[2]["disp"])


def test_super_pixel_table_and_ordered_histogram(ctx):
    """camera_info[cam].slic_to_hist_xy (floor-histogram.cpp:222-225,269-286) and the order-preserving histogram, through
    the host entry point and through the frame path, bit for bit against the sequential CPU restatement."""
    W, H, D = 384, 240, 64
    cam = synth.make_camera(W, H, D, config=2, cam=2, n_cams=4)
    sp = _sp(D, 1)
    rng = np.random.default_rng(11)
    # super-pixels of very different sizes: 8x8 cells (few points: finite sums) next to 64x64 cells (the doubling fourth
    # component overflows to inf, as it does in the reference), some unlabelled pixels, some labels beyond the table
    lab = np.where(np.arange(W)[None, :] < W // 2, (np.arange(H)[:, None] // 8) * 100 + np.arange(W)[None, :] // 8,
                   3000 + (np.arange(H)[:, None] // 64) * 10 + np.arange(W)[None, :] // 64).astype(np.int32)
    lab[rng.random((H, W)) < 0.02] = -1
    lab[5:9, 7:30] = 99999
    lab = np.ascontiguousarray(lab)
    n_labels = 3100
    still = rng.uniform(0.0, 1.0, n_labels)
    for weighting, use_still in [(api.WGT_NORM, True), (api.WGT_NONE, False), (api.WGT_QUADRANCE, True)]:
        hp = api.FloorHistogramParams(weighting=weighting, apply_slic_still_weights=use_still, ordered=True)
        desc = api.make_frame_desc(cam, sp, hp, synth.HIST_AABB, slic_lookup=lab, still_score=still)
        res = api.frame_run(ctx, 0, desc)
        out = api.frame_fetch_batch(ctx, 0, 1, W, H, res.hw, res.hh, n_labels=n_labels)[0]
        ohp = oracle.HistParams(hp.hist_sz, hp.hist_min_z, hp.hist_max_z, weighting, int(use_still))
        wh, wc = oracle.floor_hist_camera_cells(ohp, cam["et"], synth.HIST_AABB, cam["et"][:3], out["Xs"], out["xy"], lab, still)
        assert np.array_equal(out["cam_hist"], wh) and wh.sum() > 100
        got = out["slic_to_hist_xy"]
        assert got.shape == wc.shape
        assert np.array_equal(got[:, :2], wc[:, :2])
        assert np.array_equal(got[:, 2], wc[:, 2])   # inf == inf
        touched = wc[:, 0] >= 0
        assert touched.sum() > 50 and np.isinf(wc[:, 2]).any() and np.isfinite(wc[touched, 2]).any()
        assert (wc[~touched, 2] == 0).all()
    # the host entry point gives the same table
    pc = api.PointCloud()
    pc.Xs, pc.xy = out["Xs"], out["xy"]
    cells = (api._lib.SparseHistCell * n_labels)()
    hist = np.empty((res.hh, res.hw), np.float32)
    hpc = hp.to_c()
    et = (C.c_double * 8)(*cam["et"])
    ab = (C.c_double * 4)(*synth.HIST_AABB)
    cc = (C.c_double * 3)(*cam["et"][:3])
    ctx.check(ctx.lib.mvs_floor_hist_camera_cells(ctx.h, C.byref(hpc), et, ab, cc, out["Xs"].ctypes.data, out["xy"].ctypes.data, out["N"],
                                                  lab.ctypes.data, W, H, still.ctypes.data, n_labels, hist.ctypes.data, res.hw, res.hh,
                                                  cells), "mvs_floor_hist_camera_cells")
    assert np.array_equal(hist, wh)
    assert np.array_equal(np.array([(c.x, c.y, c.count) for c in cells], np.float64), wc)


def test_registered_maps_give_the_same_frame(ctx):
    """mvs_maps_register (the per-video half of get_xy_mapping): the frame path with registered maps is bit-identical."""
    import torch

    W, H, D = 384, 240, 64
    cam = synth.make_camera(W, H, D, config=2, cam=3, n_cams=4)
    sp = _sp(D, 1)
    hp = api.FloorHistogramParams(weighting=api.WGT_NONE, apply_slic_still_weights=False)
    plain = api.make_frame_desc(cam, sp, hp, synth.HIST_AABB)
    res = api.frame_run(ctx, 0, plain)
    a = api.frame_fetch(ctx, 0, W, H, res.hw, res.hh)
    handles = [api.maps_register(ctx, cam["mapx"][i], cam["mapy"][i]) for i in range(2)]
    assert handles[0] > 0 and handles[1] > 0 and handles[0] != handles[1]
    reg = api.make_frame_desc(cam, sp, hp, synth.HIST_AABB)
    for i in range(2):
        reg.maps_handle[i] = handles[i]
        reg.mapx[i] = reg.mapy[i] = None
    res2 = api.frame_run(ctx, 1, reg)
    b = api.frame_fetch(ctx, 1, W, H, res2.hw, res2.hh)
    n = W * H * 3
    dev = torch.device("cuda", ctx.device)
    for i in range(2):   # the rectified images themselves
        r0 = torch.as_tensor(bench_cai(res.rectified_dev[i], n), device=dev).cpu().numpy()
        r1 = torch.as_tensor(bench_cai(res2.rectified_dev[i], n), device=dev).cpu().numpy()
        assert np.array_equal(r0, r1)
        assert np.array_equal(r0.reshape(H, W, 3), oracle.remap_cubic_bgr8(cam["sensor"][i], cam["mapx"][i], cam["mapy"][i]))
    assert np.array_equal(a["disp"], b["disp"]) and a["N"] == b["N"] and np.array_equal(a["cam_hist"], b["cam_hist"])
    # a handle of another size, or with another interpolation, is refused
    reg.interp = api.INTER_LINEAR
    assert ctx.lib.mvs_frame_run(ctx.h, 1, C.byref(reg), None) == -1
    for hnd in handles:
        assert ctx.lib.mvs_maps_release(ctx.h, hnd) == 0
    assert ctx.lib.mvs_maps_release(ctx.h, handles[0]) == -1


class bench_cai:
    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"data": (int(ptr), False), "shape": (n,), "typestr": "|u1", "version": 2}


@pytest.mark.parametrize("params", [
    dict(block=3, p1=8, p2=32, ur=0, d12=1, sw=0, sr=0),
    dict(block=5, p1=8, p2=32, ur=10, d12=2, sw=0, sr=0),
    dict(block=9, p1=22, p2=1000, ur=3, d12=596, sw=0, sr=0),
    dict(block=13, p1=22, p2=1000, ur=25, d12=1, sw=50, sr=2),
    dict(block=11, p1=22, p2=4000, ur=3, d12=596, sw=0, sr=0),
    dict(block=5, p1=10, p2=120, ur=5, d12=1, sw=20, sr=1, minD=3),
    dict(block=7, p1=0, p2=0, ur=-1, d12=0, sw=100, sr=1, cap=10),
])
def test_sgbm_older_recurrence_form(ctx, params):
    """mvs_set_sgbm_recurrence(1): L = C + min(..) - (min Lp + P2), the form of older OpenCV sources (the reference pins
    4.4.0), against the restatement built with -DORC_SGBM_SUBTRACT_P2 (not pinned by a runnable OpenCV: cv2 here is 4.13)."""
    differs = exact = 0
    try:
        assert ctx.lib.mvs_set_sgbm_recurrence(ctx.h, 1) == 0
        for mode in (1, 0, 3):
            L, R = _pair(160, 56, 32, 5 + mode, False)
            sp = _sp(32, mode, **params)
            want = oracle.sgbm_compute(_op(sp), L, R, subtract_p2=True)
            got = np.empty((56, 160), np.int16)
            rc = ctx.lib.mvs_sgbm_compute16(ctx.h, ctypes.byref(sp.to_c()), L.ctypes.data, R.ctypes.data, 160, 56, 160, got.ctypes.data, 320)
            if rc == -3 and b"saturated" in ctx.lib.mvs_last_error(ctx.h):
                continue   # a saturated sum took part in a decision (large blocks): reported, no result claimed
            assert rc == 0
            exact += 1
            assert np.array_equal(got, want), f"mode {mode}: {(got != want).sum()} px differ"
            differs += int((want != oracle.sgbm_compute(_op(sp), L, R)).sum())
        if params["block"] <= 9:
            assert exact == 3
        # an uncorrelated pair saturates cost sums: that is reported, not silently approximated
        L, R = _pair(160, 56, 32, 9, True)
        sp = _sp(32, 1, **params)
        d16 = np.empty((56, 160), np.int16)
        rc = ctx.lib.mvs_sgbm_compute16(ctx.h, ctypes.byref(sp.to_c()), L.ctypes.data, R.ctypes.data, 160, 56, 160, d16.ctypes.data, 320)
        if rc == 0:
            assert np.array_equal(d16, oracle.sgbm_compute(_op(sp), L, R, subtract_p2=True))
        else:
            assert rc == -3 and b"saturated" in ctx.lib.mvs_last_error(ctx.h)
    finally:
        assert ctx.lib.mvs_set_sgbm_recurrence(ctx.h, 0) == 0
    if params["ur"] > 0 and exact:
        assert differs > 0   # the two forms do disagree (uniqueness test) on these inputs: the switch is not a no-op


@pytest.mark.parametrize("params", [
    dict(block=3, p1=8, p2=32, ur=10, d12=1, sw=0, sr=0),
    dict(block=7, p1=22, p2=1000, ur=3, d12=596, sw=0, sr=0),
    dict(block=13, p1=22, p2=1000, ur=25, d12=1, sw=50, sr=2),
])
@pytest.mark.parametrize("D", [256, 208])
def test_sgbm_older_recurrence_form_in_sweep(ctx, params, D):
    """The same switch at 256 disparities (and 208, padded to 256), where the winner-take-all step runs inside the backward
    sweep: its uniqueness test and its report of a saturated sum are the ones exercised here."""
    W, H = D + 150, 40
    exact = 0
    try:
        assert ctx.lib.mvs_set_sgbm_recurrence(ctx.h, 1) == 0
        for mode in (1, 0, 3):
            for noise in (False, True):
                L, R = _pair(W, H, D, 11 + mode, noise)
                sp = _sp(D, mode, **params)
                got = np.empty((H, W), np.int16)
                rc = ctx.lib.mvs_sgbm_compute16(ctx.h, ctypes.byref(sp.to_c()), L.ctypes.data, R.ctypes.data, W, H, W, got.ctypes.data, 2 * W)
                if rc == -3 and b"saturated" in ctx.lib.mvs_last_error(ctx.h):
                    continue   # reported, no result claimed
                assert rc == 0
                exact += 1
                want = oracle.sgbm_compute(_op(sp), L, R, subtract_p2=True)
                assert np.array_equal(got, want), f"mode {mode}: {(got != want).sum()} px differ"
        if params["block"] <= 7:
            assert exact >= 3
    finally:
        assert ctx.lib.mvs_set_sgbm_recurrence(ctx.h, 0) == 0


@pytest.mark.parametrize("W,H,D,dtype", [(192, 96, 64, np.uint8), (320, 120, 128, np.uint16), (384, 130, 256, np.uint8),
                                         (330, 98, 256, np.uint16)])
def test_census_sgm_strategy(ctx, W, H, D, dtype):
    """The CUDA_SGM strategy (sgm_disp.cpp:102-211 over libSGM's kernels) against its CPU restatement (parity unpinned)."""
    l, r = synth.rectified_pair(W, H, min(D, 96), 31 + D)
    gl, gr = oracle.bgr2gray(l), oracle.bgr2gray(r)
    if dtype == np.uint16:
        gl, gr = gl.astype(np.uint16) * 257, gr.astype(np.uint16) * 257
    gl[5:9, 40:60] = 0   # source pixels that are 0 mask the output (check_consistency.cu:29-31)
    want = oracle.census_sgm(gl, gr, D, taps=True)
    got, d16 = api.SGMDisparity.calculate(D, True, gl, gr, ctx=ctx, tap=True)
    assert np.array_equal(d16, want["disp16"]), f"{(d16 != want['disp16']).sum()} px differ"
    assert np.array_equal(got, want["disp"])
    assert (got[5:9, 40:60] == 0).all() and (got > 0).mean() > 0.5
