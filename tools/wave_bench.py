#!/usr/bin/env python
"""Developer tool: time the SGM stages of the batched frame path with a given build of the library.

    python tools/wave_bench.py [path/to/libmvstereo_variant.so] [--workload c2] [--batch 16] [--mode 1] [--reps 5]

Prints one JSON line with the per-frame stage times (CUDA events inside the library).  Used to compare tuning builds
(tools/build_variant.sh) on the GPU box in one gpurun call; not part of the product or of the tests.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("lib", nargs="?", default=None)
    ap.add_argument("--workload", default="c2")
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--mode", type=int, default=1)
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    from multiview_b200 import _lib
    if a.lib:
        _lib.LIB_PATH = os.path.abspath(a.lib)
    import numpy as np
    import torch

    import bench
    from multiview_b200 import api, synth

    W, H, D = bench.WORKLOADS[a.workload]
    dev = torch.device("cuda", 0)
    ctx = api.Context(0)
    sp = api.StereoSGBMParams(sg_num_disparities=D, sg_mode=a.mode)
    hp = api.FloorHistogramParams(weighting=api.WGT_NONE, apply_slic_still_weights=False)
    cams = [synth.make_camera(W, H, D, config=2, cam=c, frame=0, n_cams=2) for c in range(2)]
    keep, descs = [], []
    for cam in cams:
        dt = {"sensor": [torch.from_numpy(s).to(dev) for s in cam["sensor"]],
              "mapx": [torch.from_numpy(m).to(dev) for m in cam["mapx"]], "mapy": [torch.from_numpy(m).to(dev) for m in cam["mapy"]]}
        keep.append(dt)
        ptrs = {k: [t.data_ptr() for t in v] for k, v in dt.items()}
        descs.append(api.make_frame_desc(cam, sp, hp, synth.HIST_AABB, on_device=True, dev_ptrs=ptrs))
    B = a.batch
    batch = (api.FrameDesc * B)(*[descs[b % 2] for b in range(B)])
    stages = ["cost", "aggregate", "wta"]
    acc = {s: [] for s in stages}
    for i in range(a.reps + 2):
        api.frame_run_batch(ctx, 0, batch)
        ctx.sync()
        if i >= 2:
            for s in stages:
                acc[s].append(ctx.stage_ms(s))
    cells = bench.sgbm_cells(W, H, D)
    agg = float(np.median(acc["aggregate"])) / B
    out = {"lib": os.path.basename(_lib.LIB_PATH), "workload": a.workload, "batch": B, "mode": a.mode,
           "ms_per_frame": {s: round(float(np.median(v)) / B, 4) for s, v in acc.items()},
           "agg_frac_of_peak": round(8.0 * cells / (agg / 1000.0) / 1e9 / bench.peaks()[0], 4)}
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
