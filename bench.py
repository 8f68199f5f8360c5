#!/usr/bin/env python
"""bench.py -- stereo frames/s of the hot path (rectify + SGM + cloud + floor_hist) on N B200s.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --gpus N --steps K --warmup W   # reference CPU path (cv2 + C port)

A step is one pass of the hot path over one batch of synthetic camera-frames (default: the
BASELINE.json configs[1] shape, 1920x1080, 256 disparities, MODE_HH 8-path).  Work shards by camera:
every rank processes its own cameras, the per-step floor histograms are summed with NCCL.
Rank 0 prints ONE JSON line.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOADS = {  # name -> (W, H, D); c1..c4 are BASELINE.json configs[0..3], the rest the corners of the configs[4] sweep
    "c1": (896, 672, 128),
    "c2": (1920, 1080, 256),
    "c3": (1920, 1080, 256),    # 8-camera scene, ONE camera per GPU; a step is --batch consecutive frames of that camera
    "c4": (896, 672, 128),      # 16 cameras x 300 frames, FIXED total work sharded by camera/frame (strong scaling)
    "vga64": (640, 480, 64),
    "720p128": (1280, 720, 128),
    "1440p256": (2560, 1440, 256),
    "4k256": (3840, 2160, 256),
    "1080p192": (1920, 1080, 192),   # not a BASELINE config: the one disparity width (Dp = 192) the others do not exercise
}
C4_CAMS, C4_FRAMES, C4_FRAMES_PER_STEP = 16, 300, 2   # a c4 step = 2 frame times of all 16 cameras = 32 camera-frames, whatever N is
METRIC = "stereo frames/s (rectify+SGM+cloud+floor_hist)"


def config_dict(args, world):
    """The `config` object of the JSON line: the same for both arms (the driver compares them)."""
    W, H, D = WORKLOADS[args.workload]
    mode_name = {1: "8-path MODE_HH", 3: "4-path MODE_HH4", 0: "5-path MODE_SGBM"}[args.mode]
    cells = sgbm_cells(W, H, D)
    if args.workload == "c4":
        what = (f"{C4_CAMS} cameras x {C4_FRAMES} frames, a step = {C4_FRAMES_PER_STEP} frame times of all cameras = "
                f"{C4_CAMS * C4_FRAMES_PER_STEP} camera-frames sharded camera-major over the GPUs (fixed total work)")
    elif args.workload == "c3":
        what = f"one camera per GPU, a step = {args.batch} consecutive frames of it"
    else:
        what = f"{args.batch} camera-frames per GPU per step"
    return {"workload": f"{args.workload}: {W}x{H} stereo pair, {D} disparities, mode {args.mode} ({mode_name}), "
                        f"full disp_map+cloud+floor_hist; {what}",
            "mode": args.mode, "batch_per_gpu": units_per_rank(args, world),
            "parallelism": f"camera/frame-sharded x{world}, NCCL all-reduce of floor histograms",
            "l2": f"no flush: each cost volume is {cells * 2 / 1e6:.0f} MB > 126 MB L2",
            "distinct_inputs_per_gpu": min(args.distinct, units_per_rank(args, world))}


def units_per_rank(args, world):
    if args.workload == "c4":
        return max(1, C4_CAMS * C4_FRAMES_PER_STEP // world)
    return args.batch


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def __enter__(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None
        return self

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def __exit__(self, *a):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except Exception:
                self.proc.kill()

    def summary(self):
        sm = [float(r[0]) for r in self.rows if len(r) >= 6 and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) >= 6 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


class _CAI:
    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"data": (int(ptr), False), "shape": tuple(shape), "typestr": typestr, "version": 2}


def sgbm_cells(W, H, D, minD=1):
    return H * (W - minD - D) * D


# ------------------------------------------------------------------------------------------------ reference arm
def cpu_frame(cam, D, mode, aabb, threads=None):
    """The reference's CPU path for one camera-frame: cv2 (its OpenCV calls) for a1-a4, the C port for a6-a10."""
    import oracle
    from oracle import cv2ref

    rect = [cv2ref.remap_cubic_bgr8(cam["sensor"][i], cam["mapx"][i], cam["mapy"][i]) for i in range(2)]
    p = oracle.SgbmParams.reference_defaults(D, mode)
    disp = cv2ref.disparity_calculate(p, rect[0], rect[1], 2.15)
    b = oracle.bcam_init(cam["q"], cam["t"], cam["baseline"], cam["K"])
    pc = oracle.point_cloud(b, disp)
    hp = oracle.HistParams(0.10, 0.5, 2.0, 0, 0)
    hist = oracle.floor_hist_camera(hp, cam["et"], aabb, cam["et"][:3], pc["Xs"], pc["xy"])
    return disp, pc["N"], hist


def crop_cam(cam, rows):
    c = dict(cam)
    c["sensor"] = [np.ascontiguousarray(s[:rows]) for s in cam["sensor"]]
    c["mapx"] = [np.ascontiguousarray(m[:rows]) for m in cam["mapx"]]
    c["mapy"] = [np.ascontiguousarray(m[:rows]) for m in cam["mapy"]]
    c["H"] = rows
    return c


def run_reference(args, rank, world):
    if rank != 0:
        return
    from concurrent.futures import ThreadPoolExecutor

    from multiview_b200 import synth
    from oracle import cv2ref

    W, H, D = WORKLOADS[args.workload]
    cores = os.cpu_count() or 1
    # cv2's MODE_HH / MODE_SGBM are single-threaded inside OpenCV, so the multi-core baseline is one
    # camera-frame per thread (cv2 releases the GIL), as the reference's thread pool runs cameras in parallel.
    want = args.ref_threads if args.ref_threads > 0 else cores   # default: every host core
    par = max(1, min(cores, want, int(psutil_avail_gb() // (6.0 * W * H * D / (1920 * 1080 * 256) + 0.5)) or 1))
    cv2ref.set_threads(1 if par > 1 else cores)
    cams = [synth.make_camera(W, H, D, config=2, cam=i % 2, frame=0, n_cams=2) for i in range(min(par, 2))]
    rows = H
    t0 = time.perf_counter()
    cpu_frame(crop_cam(cams[0], min(H, 160)), D, args.mode, synth.HIST_AABB)
    est = (time.perf_counter() - t0) * H / min(H, 160)
    budget = 200.0 / max(args.steps + args.warmup, 1)
    if est * 1.3 > budget:
        rows = int(max(128, min(H, H * budget / (est * 1.3)))) // 8 * 8
    work = [crop_cam(cams[i % len(cams)], rows) for i in range(par)]

    def step():
        with ThreadPoolExecutor(par) as ex:
            list(ex.map(lambda c: cpu_frame(c, D, args.mode, synth.HIST_AABB), work))

    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    frames = par * args.steps * rows / H
    val = frames / dt
    sample = (f"{par} camera-frame(s) per step in parallel threads, each {W}x{rows} of {W}x{H} rows (throughput scaled by "
              f"rows/H), D={D}, mode={args.mode}; cv2 {cv2ref._cv2().__version__} remap+cvtColor+2xGaussianBlur+StereoSGBM "
              f"(the reference's OpenCV calls, IPP off) + C port of PointCloud::init / floor_hist")
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": "frames/s", "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1000 * dt / args.steps, "higher_is_better": True,
            "scaling": "strong" if args.workload == "c4" else "weak",
            "vs_baseline": None, "dtype": "u16", "data": "synthetic",
            "config": config_dict(args, args.gpus),
            "gpix_disp_per_s": val * W * H * D / 1e9,
            "cpu_baseline": {"value": val, "unit": "frames/s", "cores": par, "kind": "cv2+port", "sample": sample,
                             "host_cores": cores},
            "e2e": {"value": val, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def psutil_avail_gb():
    try:
        import psutil

        return psutil.virtual_memory().available / 2**30
    except Exception:
        return 16.0


# ------------------------------------------------------------------------------------------------ CUDA arm
def run_ours(args, rank, world, local_rank):
    import torch
    import torch.distributed as dist

    from multiview_b200 import api, shard, synth

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    W, H, D = WORKLOADS[args.workload]
    B = units_per_rank(args, world)
    # keep the host threads of a rank on their own share of the cores (8 ranks share one host)
    try:
        cores = sorted(os.sched_getaffinity(0))
        per = max(1, len(cores) // max(world, 1))
        mine = cores[local_rank * per:(local_rank + 1) * per] if world > 1 else cores
        if mine:
            os.sched_setaffinity(0, mine)
    except (AttributeError, OSError):
        pass
    ctx = api.Context(local_rank)
    stream = torch.cuda.ExternalStream(ctx.lib.mvs_stream(ctx.h), device=dev)
    # the library finishes a batch (WTA .. histogram, histogram sums) on its tail stream, overlapped with the next batch's
    # aggregation on the main stream; the all-reduce of the histogram sum goes behind it on that stream
    tail = torch.cuda.ExternalStream(ctx.lib.mvs_tail_stream(ctx.h), device=dev)
    sp = api.StereoSGBMParams(sg_num_disparities=D, sg_mode=args.mode)
    hp = api.FloorHistogramParams(weighting=api.WGT_NONE, apply_slic_still_weights=False)
    # the (camera, frame) units of one step that this rank owns (camera c lives on rank c % world)
    if args.workload == "c4":
        n_cams = C4_CAMS
        units = shard.units_for_rank(C4_CAMS, C4_FRAMES_PER_STEP, world, rank)
    elif args.workload == "c3":
        n_cams = world
        units = [(rank, f) for f in range(B)]
    else:
        n_cams = world * B
        units = [(c, 0) for c in shard.cameras_for_rank(n_cams, world, rank)]
    assert len(units) == B, (len(units), B)
    cfg_id = {"c1": 1, "c2": 2, "c3": 3, "c4": 4}.get(args.workload, 5)
    cams = [synth.make_camera(W, H, D, config=cfg_id, cam=c, frame=f, n_cams=n_cams) for (c, f) in units[:min(B, args.distinct)]]

    # pinned host copies (e2e) and device-resident copies (value)
    host, devt, descs_dev, descs_host = [], [], [], []
    for cam in cams:
        hc = dict(cam)
        hc["sensor"] = [torch.from_numpy(s).pin_memory() for s in cam["sensor"]]
        dt = {"sensor": [t.to(dev) for t in hc["sensor"]],
              "mapx": [torch.from_numpy(m).to(dev) for m in cam["mapx"]], "mapy": [torch.from_numpy(m).to(dev) for m in cam["mapy"]]}
        host.append(hc)
        devt.append(dt)
        ptrs = {k: [t.data_ptr() for t in v] for k, v in dt.items()}
        d = api.make_frame_desc(cam, sp, hp, synth.HIST_AABB, on_device=True, dev_ptrs=ptrs)
        if not args.float_maps:
            # rectification maps are per-video constants (get_xy_mapping.cpp:59): registered once in the fixed-point form
            # the cubic remap reads (mvs_maps_register)
            for i in range(2):
                d.maps_handle[i] = api.maps_register(ctx, cam["mapx"][i], cam["mapy"][i])
        descs_dev.append(d)
    torch.cuda.synchronize()
    hw, hh = api.hist_dims(synth.HIST_AABB, hp.hist_sz)

    def hist_reduce(first_slot=0):
        ptr, w_, h_ = C.c_void_p(), C.c_int(), C.c_int()
        ctx.check(ctx.lib.mvs_frame_hist_sum(ctx.h, first_slot, B, C.byref(ptr), C.byref(w_), C.byref(h_)), "mvs_frame_hist_sum")
        t = torch.as_tensor(_CAI(ptr.value, (hh, hw), "<f4"), device=dev)
        if world > 1:
            with torch.cuda.stream(tail):
                dist.all_reduce(t)  # counts: exact in fp32 in any order
        # floor-histogram.cpp:411-434 on the summed histogram: smoothing + clip, on the same stream
        ctx.check(ctx.lib.mvs_hist_post_dev(ctx.h, ptr, hw, hh, C.c_float(hp.gaussian_sigma), C.c_float(hp.min_hist_count),
                                            C.c_float(hp.max_hist_count)), "mvs_hist_post_dev")
        return t

    batch_dev = (api.FrameDesc * B)(*[descs_dev[b % len(descs_dev)] for b in range(B)])

    def step_dev():
        api.frame_run_batch(ctx, 0, batch_dev)  # one step = the B camera-frames of this rank, one batched call
        return hist_reduce()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- value: inputs resident in HBM --------------------------------------------------------------------
    for _ in range(args.warmup):
        step_dev()
    barrier()
    l0 = ctx.launch_count()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local_rank) as clk:
        e0.record(stream)
        for _ in range(args.steps):
            hist_t = step_dev()
        e1.record(tail)   # the last work of the last step is the histogram post-step on the tail stream
        barrier()
    ms = e0.elapsed_time(e1)
    launches = ctx.launch_count() - l0
    tmax = torch.tensor([ms], device=dev)
    if world > 1:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    ms = float(tmax.item())
    frames = (C4_CAMS * C4_FRAMES_PER_STEP if args.workload == "c4" else world * B) * args.steps
    value = frames / (ms / 1000.0)
    hist_total = float(hist_t.sum().item())

    # ---- per-stage device times (CUDA events inside the library, on the launching stream) -----------------
    # (a) the batched step, as timed above: stage time / B = per camera-frame; the aggregation stage is 2 launches
    #     (forward + backward sweep) covering all B frames, so this is its per-launch duration inside the step
    stages = ["rectify", "presmooth", "cost", "aggregate", "aggregate_fwd", "aggregate_bwd", "wta", "cloud", "hist"]
    acc = {s_: [] for s_ in stages}
    for _ in range(max(3, min(args.steps, 10))):
        api.frame_run_batch(ctx, 0, batch_dev)
        ctx.sync()
        for s_ in stages:
            acc[s_].append(ctx.stage_ms(s_))
    stage_ms = {s_: float(np.mean(v)) / B for s_, v in acc.items()}
    stage_ms["lr_check+post"] = stage_ms.pop("wta")   # the winner-take-all step itself runs inside the backward sweep
    sweep_ms = {"forward": stage_ms.pop("aggregate_fwd"), "backward_with_wta": stage_ms.pop("aggregate_bwd")}
    if not args.float_maps:   # the rectify stage when it converts the float maps in every frame
        batch_f = (api.FrameDesc * B)(*[descs_dev[b % len(descs_dev)] for b in range(B)])
        for b in range(B):
            batch_f[b].maps_handle[0] = batch_f[b].maps_handle[1] = 0
        racc = []
        for it in range(4):
            api.frame_run_batch(ctx, 0, batch_f)
            ctx.sync()
            if it:
                racc.append(ctx.stage_ms("rectify"))
        stage_ms["rectify_float_maps"] = float(np.mean(racc)) / B
    # (b) one camera-frame alone (latency mode); the first two calls load the single-frame kernels and are not counted
    acc1 = {s_: [] for s_ in stages + ["post"] if not s_.startswith("aggregate_")}
    for it in range(5):
        api.frame_run(ctx, 0, descs_dev[0])
        ctx.sync()
        if it < 2:
            continue
        for s_ in acc1:
            acc1[s_].append(ctx.stage_ms(s_))
    stage_ms_single = {s_: float(np.mean(v)) for s_, v in acc1.items()}
    # (c) the histogram stage in the reference's DEFAULT mode (WGT_NORM, SLIC still weights, floor-histogram.hpp:34-35),
    #     summed in point order (bit-exact): a stable sort by bin per camera-frame; synthetic 32x32-pixel label cells
    gx, gy = (W + 31) // 32, (H + 31) // 32
    lab = ((np.arange(H)[:, None] // 32) * gx + (np.arange(W)[None, :] // 32)).astype(np.int32)
    still = np.random.default_rng(5).uniform(0.0, 1.0, gx * gy)
    lab_t, still_t = torch.from_numpy(lab).to(dev), torch.from_numpy(still).to(dev)
    hp_ref = api.FloorHistogramParams(weighting=api.WGT_NORM, apply_slic_still_weights=True, ordered=True)
    batch_ref = (api.FrameDesc * B)(*[descs_dev[b % len(descs_dev)] for b in range(B)])
    for b in range(B):
        batch_ref[b].hist = hp_ref.to_c()
        batch_ref[b].labels, batch_ref[b].lstep = lab_t.data_ptr(), W * 4
        batch_ref[b].still_score, batch_ref[b].n_labels, batch_ref[b].labels_on_device = still_t.data_ptr(), int(still.size), 1
    hacc = []
    for it in range(4):
        api.frame_run_batch(ctx, 0, batch_ref)
        ctx.sync()
        if it:
            hacc.append(ctx.stage_ms("hist"))
    stage_ms["hist_reference_default_ordered"] = float(np.mean(hacc)) / B
    cells = sgbm_cells(W, H, D)
    peak, peak_src = peaks()
    agg_bytes = 8.0 * cells  # read C + write S forward, read C + read S backward
    achieved = agg_bytes / (stage_ms["aggregate"] / 1000.0) / 1e9
    roofline = {"bound": "hbm", "kernel": f"wave_kernel: SGM aggregation, forward + backward sweep over the batch of {B} frames; the backward "
                                          "sweep includes the winner-take-all step (SURVEY.md 8d counts no bytes for it: S stays in registers)",
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                "algorithmic_bytes_per_frame": agg_bytes, "ms_per_frame": stage_ms["aggregate"], "peak_source": peak_src,
                "frac_of_8TBs_nominal": achieved / 8000.0,
                "sweeps": {k_: {"ms_per_frame": v_, "frac": 4.0 * cells / (v_ / 1000.0) / 1e9 / peak} for k_, v_ in sweep_ms.items()},
                "single_frame": {"ms": stage_ms_single["aggregate"],
                                 "frac": agg_bytes / (stage_ms_single["aggregate"] / 1000.0) / 1e9 / peak}}

    # DRAM traffic of the same two launches from the committed ncu --set full capture (per camera-frame, like
    # algorithmic_bytes_per_frame); only quoted when the capture is of this workload
    tpath = os.path.join(ROOT, "profiles", "r02_wave_traffic.json")
    if os.path.exists(tpath):
        tj = json.load(open(tpath))
        if tj.get("workload") == args.workload and args.mode == 1:
            roofline["traffic"] = tj["traffic_bytes_per_frame"]
            roofline["traffic_source"] = "profiles/r02_wave_traffic.json (ncu dram__bytes_read+write, batch %d)" % tj["batch"]

    # ---- e2e: host buffers in, host results out, through the C ABI ----------------------------------------------
    descs_host = []
    for hc in host:
        d = api.make_frame_desc({**hc, "sensor": [t.numpy() for t in hc["sensor"]]}, sp, hp, synth.HIST_AABB)
        descs_host.append(d)
    # rectification maps are per-camera constants (get_xy_mapping is computed once per video and cached by the DAG,
    # get_xy_mapping.cpp:59), so they stay on the device; sensor images are the per-frame input.
    for b, d in enumerate(descs_host):
        for i in range(2):
            d.mapx[i] = devt[b]["mapx"][i].data_ptr()
            d.mapy[i] = devt[b]["mapy"][i].data_ptr()
            d.maps_handle[i] = descs_dev[b].maps_handle[i]
    # two slot ranges alternate: while batch k's kernels run, batch k+1's sensor images upload and batch k-1's
    # results download on the library's copy stream
    from multiview_b200._lib import FetchDesc

    out_disp = [[torch.empty((H, W), dtype=torch.float32).pin_memory() for _ in range(B)] for _ in range(2)]
    out_hist = [[torch.empty((hh, hw), dtype=torch.float32).pin_memory() for _ in range(B)] for _ in range(2)]
    out_cen = [np.empty((B, 3)) for _ in range(2)]
    nouts = [[C.c_int32() for _ in range(B)] for _ in range(2)]
    batch_host = (api.FrameDesc * B)(*[descs_host[b % len(descs_host)] for b in range(B)])
    for b in range(B):
        batch_host[b].inputs_on_device, batch_host[b].maps_on_device = 0, 1  # host (pinned) sensor images, resident maps

    def make_fetch(with_cloud):
        # one descriptor table per slot range; with_cloud adds the host PointCloud (Xs, xy, lookup) a consumer of
        # disp_map_update::Result receives (pipeline/nodes/disp_map_update.cpp:115-119)
        tabs, keep = [], []
        for half in range(2):
            t = (FetchDesc * B)()
            for b in range(B):
                t[b].disp, t[b].dstep = out_disp[half][b].data_ptr(), W * 4
                t[b].cam_hist = out_hist[half][b].data_ptr()
                t[b].centroid = out_cen[half][b].ctypes.data
                t[b].n_out = C.pointer(nouts[half][b])
                if with_cloud:
                    xs = torch.empty((H * W, 3), dtype=torch.float64).pin_memory()
                    xy = torch.empty((H * W, 2), dtype=torch.int32).pin_memory()
                    lk = torch.empty((H, W), dtype=torch.int32).pin_memory()
                    keep.append((xs, xy, lk))
                    t[b].Xs, t[b].xy, t[b].lookup = xs.data_ptr(), xy.data_ptr(), lk.data_ptr()
            tabs.append(t)
        return tabs, keep

    def run_e2e(n, tabs):
        # two slot ranges alternate: batch k+1 is enqueued (uploads on the copy stream) before batch k is fetched
        pending = []
        for k in range(n):
            first = (k & 1) * B
            api.frame_run_batch(ctx, first, batch_host)
            hist_reduce(first)
            pending.append(k)
            if len(pending) > 1:
                j = pending.pop(0)
                ctx.check(ctx.lib.mvs_frame_fetch_batch(ctx.h, (j & 1) * B, B, tabs[j & 1]), "mvs_frame_fetch_batch")
        while pending:
            j = pending.pop(0)
            ctx.check(ctx.lib.mvs_frame_fetch_batch(ctx.h, (j & 1) * B, B, tabs[j & 1]), "mvs_frame_fetch_batch")
        torch.cuda.synchronize()

    def time_e2e(tabs):
        run_e2e(min(args.warmup, 3), tabs)
        barrier()
        t0 = time.perf_counter()
        run_e2e(args.steps, tabs)
        barrier()
        t2 = torch.tensor([time.perf_counter() - t0], device=dev)
        if world > 1:
            dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        return frames / float(t2.item())

    tabs, _ = make_fetch(False)
    e2e_val = time_e2e(tabs)
    h2d = B * 2 * H * W * 3
    d2h = B * (H * W * 4 + hh * hw * 4 + 4 + 24)
    e2e_cloud = None
    if not args.no_cloud_e2e:
        tabs_c, keep_c = make_fetch(True)
        e2e_cloud_val = time_e2e(tabs_c)
        n_pts = int(np.mean([v.value for v in nouts[0]]))
        e2e_cloud = {"value": e2e_cloud_val, "unit": "frames/s", "d2h_bytes_per_step": d2h + B * (n_pts * 32 + H * W * 4),
                     "note": "as e2e, plus the host PointCloud of every frame (Xs, xy, lookup), copied by exact size N"}
        del keep_c

    line = {"metric": METRIC, "value": value, "unit": "frames/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "strong" if args.workload == "c4" else "weak",
            "vs_baseline": None, "dtype": "u16", "data": "synthetic",
            "config": config_dict(args, world),
            "gpix_disp_per_s": value * W * H * D / 1e9,
            "roofline": roofline, "stage_ms": stage_ms, "stage_ms_single_frame": stage_ms_single, "gpu_launches": int(launches),
            "e2e": {"value": e2e_val, "unit": "frames/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "note": "sensor images H2D from pinned memory + float disparity, floor histogram, N, centroid D2H every frame; "
                            "the point cloud stays on the device (its consumer floor_hist runs there)"},
            "clocks": clk.summary(), "hist_mass_after_smoothing": hist_total}
    if e2e_cloud:
        line["e2e_with_cloud"] = e2e_cloud

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        from oracle import cv2ref

        cores = os.cpu_count() or 1
        cv2ref.set_threads(cores)
        rows = min(H, args.cpu_rows)
        c = crop_cam(cams[0], rows)
        t0 = time.perf_counter()
        cpu_frame(c, D, args.mode, synth.HIST_AABB)
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": (rows / H) / dt, "unit": "frames/s", "cores": cores, "kind": "cv2+port",
                                "sample": f"one camera-frame, first {rows} of {H} rows ({W}x{rows}, D={D}, mode {args.mode}), scaled by rows/H; "
                                          f"cv2 {cv2ref._cv2().__version__} with setNumThreads({cores}) (StereoSGBM MODE_HH itself is "
                                          f"single-threaded inside OpenCV) + C port of cloud/hist; `--impl reference` runs one frame per core",
                                "seconds": dt}
    if rank == 0:
        emit(line)
    if world > 1:
        dist.destroy_process_group()


_RESULT_FD = None


def _claim_stdout():
    """stdout carries exactly ONE JSON line: libraries that write there on their own (NCCL prints its version banner to
    stdout on communicator creation) are sent to stderr at the file-descriptor level; emit() uses the saved descriptor."""
    global _RESULT_FD
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line: dict):
    data = (json.dumps(line) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_RESULT_FD, data)


def main():
    _claim_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2", choices=sorted(WORKLOADS))
    ap.add_argument("--mode", type=int, default=1, help="StereoSGBM mode: 1 = MODE_HH (8-path, headline), 3 = HH4, 0 = SGBM")
    ap.add_argument("--batch", type=int, default=16, help="camera-frames per GPU per step (<= 32; 32 measured no faster than 16)")
    ap.add_argument("--distinct", type=int, default=4, help="distinct synthetic camera-frames generated per rank")
    ap.add_argument("--ref-threads", type=int, default=0, help="reference arm: camera-frames in parallel (0 = every host core)")
    ap.add_argument("--float-maps", action="store_true", help="do not register the rectification maps (convert them in every frame)")
    ap.add_argument("--no-cloud-e2e", action="store_true", help="skip the second end-to-end measurement (point clouds to the host)")
    ap.add_argument("--cpu-rows", type=int, default=1080)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
