"""Builds csrc/libgsr.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m gaussiansplattingviewer_b200.build [--force] [--verbose]

preprocess.cu is compiled with -fmad=false (bit-exact records, see its header); everything else with
the default FMA contraction.  -lineinfo is always on so ncu's source page maps to these files.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(CSRC, "libgsr.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden"]
SOURCES = {
    "preprocess.cu": ["-fmad=false"],
    "radix_sort.cu": [],
    "binning.cu": [],
    "blend.cu": [],
    "instrument.cu": [],
    "ply_decode.cu": ["-fmad=false"],
    "gsr_api.cu": ["-Xcompiler", "-fvisibility=default"],
}


def _stale(target: str, deps) -> bool:
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    hdrs = [os.path.join(CSRC, "gsr_internal.h"), os.path.join(HERE, "..", "include", "gsr.h"),
            os.path.join(HERE, "..", "include", "gsr_measure.h"), os.path.abspath(__file__)]
    objs = []
    for src, extra in SOURCES.items():
        s = os.path.join(CSRC, src)
        o = os.path.join(CSRC, src.replace(".cu", ".o"))
        objs.append(o)
        if force or _stale(o, [s] + hdrs):
            cmd = [NVCC] + ARCH + COMMON + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o]
            if verbose:
                print(" ".join(cmd), file=sys.stderr)
            subprocess.run(cmd, check=True)
    if force or _stale(LIB, objs):
        cmd = [NVCC] + ARCH + ["-shared", "-cudart", "static", "-o", LIB] + objs
        if verbose:
            print(" ".join(cmd), file=sys.stderr)
        subprocess.run(cmd, check=True)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv or "-v" in sys.argv))
