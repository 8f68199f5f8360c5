#!/usr/bin/env python
"""Developer tool (GPU box): the CUDA_SGM strategy end to end (host images in, host disparity out), this library against the
reference's vendored libSGM recompiled for sm_100a (oracle/_ref/libsgm_ref.so, `make -C oracle ref`), same inputs.  Wall clock
around the blocking calls, best of a few repetitions; prints one JSON line per shape.
"""
import ctypes
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402  (images only: bgr2gray)
from multiview_b200 import api, synth  # noqa: E402


def main():
    ref = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libsgm_ref.so"))
    ctx = api.Context(0)
    for (W, H, D) in [(1920, 1080, 128), (1280, 720, 128), (1280, 720, 64)]:
        l, r = synth.rectified_pair(W, H, min(D, 96), 5)
        gl, gr = np.ascontiguousarray(oracle.bgr2gray(l)), np.ascontiguousarray(oracle.bgr2gray(r))
        dst = np.zeros((H, W), np.uint16)

        def run_ref():
            assert ref.sgm_ref_execute(gl.ctypes.data_as(ctypes.c_void_p), gr.ctypes.data_as(ctypes.c_void_p),
                                       dst.ctypes.data_as(ctypes.c_void_p), W, H, D, 8, 16) == 0

        def run_ours():
            api.SGMDisparity.calculate(D, True, gl, gr, ctx=ctx)

        out = {"shape": [W, H, D]}
        for name, fn in (("libsgm_ref_ms", run_ref), ("ours_ms", run_ours)):
            for _ in range(3):
                fn()
            ts = []
            for _ in range(8):
                t0 = time.perf_counter()
                fn()
                ts.append((time.perf_counter() - t0) * 1e3)
            out[name] = round(min(ts), 3)
        out["note"] = ("host to host; libsgm_ref constructs sgm::StereoSGM (cudaMalloc of its buffers) in every call, as "
                       "SGMDisparity::calculate does (sgm_disp.cpp:153); ours reuses its arena")
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
