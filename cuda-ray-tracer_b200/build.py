"""Build librt_b200.so (CUDA, sm_100a) in-tree with nvcc.  Cross-compiles without a GPU.

    python cuda-ray-tracer_b200/build.py [--force]

The exact-arithmetic translation units are compiled with -fmad=false so every FP64/FP32 operation rounds
where the reference CPU path rounds (rt_math.cuh).  Kernels that want fused multiply-adds (the FP32
sphere filter) ask for them explicitly with fmaf()/__fmaf_rn.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "librt_b200.so")
SOURCES = ["rt_api.cu", "rt_kernels.cu", "rt_wave.cu", "rt_wave_plain.cu", "rt_wave_grid.cu", "rt_wave_grid_plain.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-fmad=false", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Xcompiler", "-Wall", "--threads", "7"]


def _stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "rt_b200.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, checked: bool = False) -> str:
    """checked=True: librt_b200_checked.so with -DRT_B200_CHECKED (index assertions in the kernels, rt_wave.cuh: RT_CHECK)."""
    if checked:
        out = os.path.join(HERE, "librt_b200_timeline.so" if checked == "timeline" else "librt_b200_checked.so")
        nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
        subprocess.run([nvcc] + NVCC_FLAGS + ["-DRT_B200_TIMELINE" if checked == "timeline" else "-DRT_B200_CHECKED", "-shared", "-o", out] +
                       [os.path.join(CSRC, s) for s in SOURCES] + ["-ldl"], check=True)
        return out
    if not force and not _stale():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-shared", "-o", LIB] + \
          [os.path.join(CSRC, s) for s in SOURCES] + ["-ldl"]
    subprocess.run(cmd, check=True)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, checked="timeline" if "--timeline" in sys.argv else ("--checked" in sys.argv)))
