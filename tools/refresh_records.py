#!/usr/bin/env python
"""Developer tool: copy the records of the last tools/final_measure.sh run from gpurun_out/ into profiles/, regenerate the ncu
summaries (details text, DRAM traffic record, launch-list summary) and rewrite the numbers DESIGN.md / README.md quote from them.

    python tools/refresh_records.py
"""
import csv
import io
import json
import os
import re
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
os.chdir(ROOT)
BENCH = ["r02_bench_c2_n1", "r02_bench_c1_n1", "r02_bench_c4_n1", "r02_bench_c2_batch1_n1", "r02_bench_c5_vga64_n1", "r02_bench_c5_720p128_n1",
         "r02_bench_c5_1440p256_n1", "r02_bench_c5_4k256_n1", "r02_bench_reference_cpu"]


def rec(n):
    return json.loads(open(f"profiles/{n}.json").read().strip().splitlines()[-1])


def ncu(args):
    return subprocess.run(["ncu"] + args, capture_output=True, text=True).stdout


def main():
    for f in BENCH:
        if os.path.exists(f"gpurun_out/{f}.json"):
            shutil.copy(f"gpurun_out/{f}.json", f"profiles/{f}.json")
    if os.path.exists("gpurun_out/r02_launch_list_c2_batch16.csv"):
        shutil.copy("gpurun_out/r02_launch_list_c2_batch16.csv", "profiles/r02_launch_list_c2_batch16.csv")
    if os.path.exists("gpurun_out/r02_wave_final.ncu-rep"):
        open("profiles/r02_wave_kernel_c2_batch16_ncu_details.txt", "w").write(ncu(["-i", "gpurun_out/r02_wave_final.ncu-rep", "--page", "details"]))
        rows = list(csv.reader(io.StringIO(ncu(["-i", "gpurun_out/r02_wave_final.ncu-rep", "--page", "raw", "--csv"]))))
        ix = {h: i for i, h in enumerate(rows[0])}
        scale = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0}

        def col(name):
            return [float(r[ix[name]]) for r in rows[2:]], rows[1][ix[name]]
        rd, u1 = col("dram__bytes_read.sum")
        wr, u2 = col("dram__bytes_write.sum")
        t, u3 = col("gpu__time_duration.sum")
        tot = sum(rd) * scale[u1] + sum(wr) * scale[u2]
        tj = json.load(open("profiles/r02_wave_traffic.json"))
        B = tj["batch"]
        tj.update({"dram_read_bytes_per_launch": [x * scale[u1] for x in rd], "dram_write_bytes_per_launch": [x * scale[u2] for x in wr],
                   "traffic_bytes_per_launch_pair": tot, "traffic_bytes_per_frame": tot / B,
                   "traffic_over_algorithmic": tot / B / tj["algorithmic_bytes_per_frame"],
                   "launch_ms": [x * (1e-3 if u3 == "us" else 1.0) for x in t]})
        json.dump(tj, open("profiles/r02_wave_traffic.json", "w"), indent=1)
    if os.path.exists("gpurun_out/r02_cost_final.ncu-rep"):
        open("profiles/r02_cost_kernel_c2_ncu_details.txt", "w").write(ncu(["-i", "gpurun_out/r02_cost_final.ncu-rep", "--page", "details"]))

    d = rec("r02_bench_c2_n1")
    r, st = d["roofline"], d["stage_ms"]
    tj = json.load(open("profiles/r02_wave_traffic.json"))
    ref = rec("r02_bench_reference_cpu")
    tot = st["rectify"] + st["presmooth"] + st["cost"] + st["aggregate"] + st["lr_check+post"] + st["cloud"] + st["hist"]
    fw, bw = r["sweeps"]["forward"], r["sweeps"]["backward_with_wta"]

    s = open("DESIGN.md").read()

    def resub(pat, new, text, count=1):
        out, n = re.subn(pat, new, text, count=count, flags=re.S)
        assert n == count, pat
        return out
    s = resub(r"forward sweep [\d.]+ ms per frame = \*\*[\d.]+ of the measured HBM peak\*\*",
              f"forward sweep {fw['ms_per_frame']:.3f} ms per frame = **{fw['frac']:.2f} of the measured HBM peak**", s)
    s = resub(r"winner-take-all step [\d.]+ ms = [\d.]+; both: \*\*[\d.]+ ms per frame = [\d.]+ of peak\*\*",
              f"winner-take-all step {bw['ms_per_frame']:.3f} ms = {bw['frac']:.2f}; both: **{r['ms_per_frame']:.3f} ms per frame = {r['frac']:.2f} of peak**", s)
    s = resub(r"DRAM traffic of the two launches:\n[\d.]+ GB per frame = [\d.]+ × the algorithmic",
              f"DRAM traffic of the two launches:\n{tj['traffic_bytes_per_frame'] / 1e9:.2f} GB per frame = {tj['traffic_over_algorithmic']:.2f} × the algorithmic", s)
    s = resub(r"A single frame alone takes [\d.]+ ms", f"A single frame alone takes {r['single_frame']['ms']:.1f} ms", s)
    s = resub(r"rectify [\d.]+ \+ grey/blur [\d.]+ \+ cost\n[\d.]+ \+ aggregation with WTA [\d.]+ \+ left-right check / median / speckle [\d.]+ \+ cloud [\d.]+ \+ histogram [\d.]+\n"
              r"= [\d.]+ ms of stage time → \*\*\d+ frames/s\*\* with inputs in HBM, \*\*\d+ frames/s\*\* end to end from pinned host images\n"
              r"\(\d+ with every frame's host `PointCloud` delivered as well\); the reference's CPU path on the same box runs\n[\d.]+ frames/s",
              f"rectify {st['rectify']:.3f} + grey/blur {st['presmooth']:.3f} + cost\n{st['cost']:.3f} + aggregation with WTA {st['aggregate']:.3f} + left-right check / median / speckle "
              f"{st['lr_check+post']:.3f} + cloud {st['cloud']:.3f} + histogram {st['hist']:.3f}\n= {tot:.2f} ms of stage time → **{d['value']:.0f} frames/s** with inputs in HBM, "
              f"**{d['e2e']['value']:.0f} frames/s** end to end from pinned host images\n({d['e2e_with_cloud']['value']:.0f} with every frame's host `PointCloud` delivered as well); "
              f"the reference's CPU path on the same box runs\n{ref['value']:.2f} frames/s", s)

    def o(n, label):
        x = rec(n)
        return f"{label} {x['value']:.0f} / {x['e2e']['value']:.0f} / {x['roofline']['frac']:.2f}"
    others = "; ".join([o("r02_bench_c5_vga64_n1", "640×480×64"), o("r02_bench_c5_720p128_n1", "1280×720×128"), o("r02_bench_c1_n1", "c1 896×672×128"),
                        o("r02_bench_c4_n1", "c4 (16 cameras × 300 frames of 896×672×128, 32 camera-frames per step)"),
                        o("r02_bench_c5_1440p256_n1", "2560×1440×256 (batch 8)"), o("r02_bench_c5_4k256_n1", "3840×2160×256 (batch 4)"),
                        o("r02_bench_c2_batch1_n1", "c2 one pair at a time (batch 1)")])
    i = s.find("of HBM peak; batch 16 unless noted; one record each under `profiles/r02_bench_*`):")
    j = s.find("Narrow disparity ranges leave most lanes")
    assert 0 < i < j
    s = s[:i] + "of HBM peak; batch 16 unless noted; one record each under `profiles/r02_bench_*`): " + others + ".\n" + s[j:]
    open("DESIGN.md", "w").write(s)

    t = open("README.md").read()
    t = resub(r"\d+ frames/s with inputs resident in HBM, \d+ frames/s end to end", f"{d['value']:.0f} frames/s with inputs resident in HBM, {d['e2e']['value']:.0f} frames/s end to end", t)
    t = resub(r"\(\d+ with every frame's point cloud", f"({d['e2e_with_cloud']['value']:.0f} with every frame's point cloud", t)
    t = resub(r"against [\d.]+ frames/s for the", f"against {ref['value']:.2f} frames/s for the", t)
    t = resub(r"runs at [\d.]+ of the measured HBM peak on its 8 B per\ncell \(forward sweep alone [\d.]+;",
              f"runs at {r['frac']:.2f} of the measured HBM peak on its 8 B per\ncell (forward sweep alone {fw['frac']:.2f};", t)
    t = resub(r"with [\d.]+ × the algorithmic\nDRAM traffic", f"with {tj['traffic_over_algorithmic']:.2f} × the algorithmic\nDRAM traffic", t)
    open("README.md", "w").write(t)

    # launch-list summary
    body = subprocess.run([sys.executable, "tools/launch_summary.py", "profiles/r02_launch_list_c2_batch16.csv"], capture_output=True, text=True).stdout
    total = float(re.search(r"([\d.]+) ms of kernel time", body).group(1))
    side = sum(float(m.group(2)) for line in body.splitlines()
               for m in [re.match(r"\| `(rs_\w+|hist_ordered_sum_kernel|slic_cells_kernel|seg_bounds_kernel|hist_points_kernel|gather_f\d_kernel)` \| \d+ \| ([\d.]+)", line)] if m)
    wave = sum(float(m.group(1)) for m in re.finditer(r"\| `wave_kernel[^`]*` \| \d+ \| ([\d.]+)", body))
    old = open("profiles/r02_launch_list_c2_batch16_summary.md").read()
    hdr = old[:old.find(re.search(r"\d+ launches captured", old).group(0))]
    hdr = re.sub(r"`gather_\*`: [\d.]+ ms that are", f"`gather_*`: {side:.1f} ms that are", hdr)
    hdr = re.sub(r"instantiations are [\d.]+ of [\d.]+ ms = [\d.]+ of the kernel", f"instantiations are {wave:.1f} of {total - side:.1f} ms = {wave / (total - side):.2f} of the kernel", hdr)
    hdr = re.sub(r"aggregation stage at [\d.]+ of [\d.]+ ms of stage time per frame = [\d.]+", f"aggregation stage at {st['aggregate']:.3f} of {tot:.3f} ms of stage time per frame = {st['aggregate'] / tot:.2f}", hdr)
    open("profiles/r02_launch_list_c2_batch16_summary.md", "w").write(hdr + body)
    print("c2", round(d["value"], 1), round(d["e2e"]["value"], 1), "frac", round(r["frac"], 4), "traffic", round(tj["traffic_over_algorithmic"], 3), "ref", round(ref["value"], 2))


if __name__ == "__main__":
    main()
