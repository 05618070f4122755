"""Builds libd3ga_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m d3ga_b200.build [--force] [--verbose]

The built library is git-ignored but travels to the GPU box with the repository snapshot.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libd3ga_b200.so")
OBJ = os.path.join(HERE, "build")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")

# (source, extra flags).  raster_geometry.cu decides integers (radii, rects, keys): no implicit FMA contraction there.
SOURCES = [
    ("api.cu", []),
    ("raster_geometry.cu", ["-fmad=false"]),
    ("raster_sort.cu", []),
    ("raster_direct.cu", []),
    ("raster_blend.cu", []),
    ("raster_backward.cu", []),
    ("raster_backward_noalloc.cu", []),
    ("deform.cu", []),
    ("fem.cu", []),
    ("ssim.cu", []),
    ("bary.cu", []),
    ("lbs.cu", []),
]
COMMON = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
]


def _newer(src_paths, target):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in src_paths)


def build_library(force: bool = False, verbose: bool = False) -> str:
    headers = [os.path.join(CSRC, "common.cuh"), os.path.join(HERE, "..", "include", "d3ga_b200.h")]
    os.makedirs(OBJ, exist_ok=True)
    objs, rebuilt = [], False
    for name, extra in SOURCES:
        src = os.path.join(CSRC, name)
        obj = os.path.join(OBJ, name.replace(".cu", ".o"))
        objs.append(obj)
        if force or _newer([src] + headers, obj):
            cmd = [NVCC] + COMMON + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
            if verbose:
                print(" ".join(cmd), file=sys.stderr)
            subprocess.check_call(cmd)
            rebuilt = True
    if rebuilt or not os.path.exists(OUT):
        cmd = [NVCC, "-shared", "-cudart", "static", "-gencode", "arch=compute_100a,code=sm_100a", "-o", OUT] + objs
        subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
