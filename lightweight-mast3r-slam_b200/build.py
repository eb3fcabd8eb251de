#!/usr/bin/env python
"""Compile libm3r_b200.so in-tree for sm_100a (nvcc cross-compiles without a GPU).

Two libraries come out of the same sources:
  libm3r_b200.so             the product (what mast3r_slam_backends loads)
  tests/_variants/libm3r_b200_variants.so
                             the same code with -DM3R_TEST_VARIANTS -DM3R_COOP_PROF: the measured-
                             slower alternative solvers (8-CTA cluster Cholesky, multi-launch Cholesky),
                             the solver switch M3R_GN_SOLVER and the clock64() phase profile of the
                             cooperative solver.  Only tests/ and tools/ load it (M3R_B200_LIB).
"""
import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
SRC = [os.path.join(HERE, "csrc", f) for f in ("common.cu", "matching.cu", "tracking.cu", "pointmap.cu", "gn.cu")]
OUT = os.path.join(HERE, "libm3r_b200.so")
VARIANTS_OUT = os.path.join(ROOT, "tests", "_variants", "libm3r_b200_variants.so")
NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-Xcompiler", "-fPIC",
]
VARIANT_DEFS = ["-DM3R_TEST_VARIANTS", "-DM3R_COOP_PROF"]


def _deps():
    csrc = os.path.join(HERE, "csrc")
    return [os.path.join(csrc, f) for f in os.listdir(csrc) if f.endswith((".cu", ".cuh"))] + [
        os.path.join(ROOT, "include", "m3r_backends.h"), os.path.abspath(__file__)]


def needs_build(out=OUT):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(d) > t for d in _deps())


def _compile(out, defs, verbose):
    """One object per source file (in parallel), then link."""
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    objdir = os.path.join(HERE, "build", os.path.basename(out) + ".o")
    os.makedirs(objdir, exist_ok=True)
    os.makedirs(os.path.dirname(out), exist_ok=True)

    def one(src):
        obj = os.path.join(objdir, os.path.basename(src) + ".o")
        cmd = [nvcc] + NVCC_FLAGS + defs + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, src]
        print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)
        return obj

    with ThreadPoolExecutor(max_workers=len(SRC)) as ex:
        objs = list(ex.map(one, SRC))
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", out] + objs
    print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    return out


def build(force=False, verbose=False, variants=True):
    if force or needs_build(OUT):
        _compile(OUT, [], verbose)
    if variants and (force or needs_build(VARIANTS_OUT)):
        _compile(VARIANTS_OUT, VARIANT_DEFS, verbose)
    return OUT


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv, variants="--no-variants" not in sys.argv)
