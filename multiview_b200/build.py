"""Builds libmvstereo.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m multiview_b200.build [--force]

The fp64 geometry units (cloud.cu, hist.cu) are compiled with --fmad=false and host code with
-ffp-contract=off so that every product/sum is rounded once, like the reference's baseline x86-64 build.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_obj")
LIB = os.path.join(HERE, "libmvstereo.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

UNITS = {
    "ctx.cu": [],
    "remap.cu": [],
    "presmooth.cu": [],
    "cost.cu": os.environ.get("MVS_COST_FLAGS", "").split(),
    # tuning experiments: MVS_WAVE_FLAGS="-DWAVE_R=15 -DWAVE_D12=7" python -m multiview_b200.build --force
    "aggregate_wave.cu": os.environ.get("MVS_WAVE_FLAGS", "").split(),
    "wta_post.cu": [],
    "census.cu": [],
    "sgbm.cu": [],
    "frame.cu": [],
    "cloud.cu": ["--fmad=false"],
    "hist.cu": ["--fmad=false"],
}
COMMON = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler",
          "-fPIC,-ffp-contract=off", "-Xptxas", "-v"]


def _headers():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hs.append(os.path.join(HERE, "..", "include", "mvstereo.h"))
    return hs


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _compile(unit, extra, log):
    src = os.path.join(CSRC, unit)
    obj = os.path.join(OBJ, unit.replace(".cu", ".o"))
    if not _stale(obj, [src] + _headers()):
        return obj
    cmd = [NVCC] + COMMON + extra + ["-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    log[unit] = r.stderr
    if r.returncode:
        raise RuntimeError(f"nvcc failed for {unit}:\n{r.stdout}\n{r.stderr}")
    return obj


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    if force:
        for f in os.listdir(OBJ):
            os.remove(os.path.join(OBJ, f))
    log = {}
    with ThreadPoolExecutor(max_workers=8) as ex:
        objs = list(ex.map(lambda kv: _compile(kv[0], kv[1], log), UNITS.items()))
    if verbose:
        for u, txt in log.items():
            print(f"== {u}\n{txt}")
    if force or _stale(LIB, objs):
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode:
            raise RuntimeError(f"link failed:\n{r.stdout}\n{r.stderr}")
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
