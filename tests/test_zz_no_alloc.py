"""mvs_point_cloud_static_dev: the point cloud for hosts that cannot allocate GPU memory at frame time.

Such hosts (a real-time camera loop that cannot afford cudaMalloc's device-wide stall, or an application
whose own pool already holds all GPU memory) hand the library device buffers for inputs and outputs and the
library keeps its scratch in a fixed device array.  This file checks, through the C ABI, that the entry point
gives the same cloud as mvs_point_cloud on two consecutive calls (the second with different data, so stale
scratch contents would show), and that oversize images are refused.  It is named test_zz_* so that pytest
collects it last.
"""
import ctypes as C

import numpy as np
import pytest

from multiview_b200 import api, synth

pytestmark = pytest.mark.gpu


def _disparity(w, h, seed, holes):
    rng = np.random.default_rng(seed)
    d = rng.uniform(1.5, 60.0, (h, w)).astype(np.float32)
    d[rng.random((h, w)) < holes] = 0.0
    d[0, :3] = [np.nan, np.inf, 0.0]
    return d


def _run_static(ctx, bcam_c, disp):
    import torch

    h, w = disp.shape
    dev = torch.device("cuda", ctx.device)
    d_disp = torch.from_numpy(disp).to(dev)
    # outputs pre-filled with garbage: every entry the caller may read must be written by the call
    Xs = torch.full((h * w, 3), -7.0, dtype=torch.float64, device=dev)
    xy = torch.full((h * w, 2), -7, dtype=torch.int32, device=dev)
    lookup = torch.full((h, w), -7, dtype=torch.int32, device=dev)
    n = torch.full((1,), -7, dtype=torch.int32, device=dev)
    cen = torch.full((3,), -7.0, dtype=torch.float64, device=dev)
    torch.cuda.synchronize()  # the library runs on its own stream
    rc = ctx.lib.mvs_point_cloud_static_dev(ctx.h, C.byref(bcam_c), d_disp.data_ptr(), w, h, 1, Xs.data_ptr(), xy.data_ptr(),
                                            lookup.data_ptr(), n.data_ptr(), cen.data_ptr())
    ctx.check(rc, "mvs_point_cloud_static_dev")
    ctx.sync()
    N = int(n.item())
    return {"N": N, "Xs": Xs[:N].cpu().numpy(), "xy": xy[:N].cpu().numpy(), "lookup": lookup.cpu().numpy(), "C": cen.cpu().numpy()}


def _oracle_bcam(bcam_c):
    import oracle

    ob = oracle.Bcam()
    for i in range(2):
        for k in range(9):
            ob.krr_inv[i][k] = bcam_c.krr_inv[i][k]
        for k in range(3):
            ob.c[i][k] = bcam_c.c[i][k]
    return ob


def test_static_scratch_point_cloud_matches_allocating_entry_point(ctx):
    import oracle
    import torch

    bc = api.BinocularCamera().init(synth.bcam_quaternion(), synth.BCAM_T, synth.BCAM_BASELINE, synth.intrinsics(640, 360))
    bcam_c = bc.to_c()
    # second call: smaller image, far fewer valid pixels -> leftovers of the first call sit beyond what it uses
    for call, (w, h, seed, holes) in enumerate([(640, 360, 1, 0.25), (333, 211, 2, 0.9)]):
        disp = _disparity(w, h, seed, holes)
        want = api.PointCloud().init(disp, True, bc, ctx=ctx)
        if call == 1:
            _run_static(ctx, bcam_c, disp)   # (the test's own torch buffers of this size now sit in torch's cache)
            torch.cuda.synchronize()
            free_before = torch.cuda.mem_get_info(ctx.device)[0]
        got = _run_static(ctx, bcam_c, disp)
        if call == 1:   # the library took no device memory for it: what the driver reports as free is unchanged
            assert torch.cuda.mem_get_info(ctx.device)[0] == free_before
        assert got["N"] == want.N() and got["N"] > 100
        assert np.array_equal(got["xy"], want.xy) and np.array_equal(got["lookup"], want.lookup)
        assert np.abs(got["Xs"] - want.Xs).max() <= 1e-4 * np.abs(want.Xs).max()  # tolerance of the other cloud tests
        assert np.allclose(got["C"], want.C, rtol=1e-4, atol=0)
        # and against the CPU restatement of PointCloud::init, not only against the sibling entry point
        ref = oracle.point_cloud(_oracle_bcam(bcam_c), disp)
        assert got["N"] == ref["N"] and np.array_equal(got["xy"], ref["xy"]) and np.array_equal(got["lookup"], ref["lookup"])
        assert np.abs(got["Xs"] - ref["Xs"]).max() <= 1e-4 * np.abs(ref["Xs"]).max()
        assert np.allclose(got["C"], ref["C"], rtol=1e-4, atol=0)


def test_static_scratch_at_exactly_the_limit(ctx):
    """4096 x 2304 pixels is the largest image the fixed scratch serves: every one of its tile entries is in use."""
    import oracle

    w, h = 4096, 2304
    bc = api.BinocularCamera().init(synth.bcam_quaternion(), synth.BCAM_T, synth.BCAM_BASELINE, synth.intrinsics(w, h))
    bcam_c = bc.to_c()
    disp = _disparity(w, h, 3, 0.5)
    got = _run_static(ctx, bcam_c, disp)
    ref = oracle.point_cloud(_oracle_bcam(bcam_c), disp)
    assert got["N"] == ref["N"] and got["N"] > 4_000_000
    assert np.array_equal(got["xy"], ref["xy"]) and np.array_equal(got["lookup"], ref["lookup"])
    assert np.abs(got["Xs"] - ref["Xs"]).max() <= 1e-4 * np.abs(ref["Xs"]).max()
    assert np.allclose(got["C"], ref["C"], rtol=1e-4, atol=0)


def test_static_scratch_rejects_oversize(ctx):
    import torch

    dev = torch.device("cuda", ctx.device)
    dummy = torch.zeros(16, dtype=torch.float64, device=dev)
    b = api.BinocularCamera().init(synth.bcam_quaternion(), synth.BCAM_T, synth.BCAM_BASELINE, synth.intrinsics(64, 64)).to_c()
    p = dummy.data_ptr()
    rc = ctx.lib.mvs_point_cloud_static_dev(ctx.h, C.byref(b), p, 4096, 2305, 1, p, p, p, p, p)  # one row past the limit
    assert rc == -3
    assert b"exceeds the fixed scratch" in ctx.lib.mvs_last_error(ctx.h)
