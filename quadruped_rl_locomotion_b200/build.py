"""Builds the CUDA library in-tree with nvcc for sm_100a (no JIT cache: the .so travels with the repo).

    libgo1sim.so      fp32 product build
    libgo1sim_f64.so  fp64 check build (same source, -DGO1_REAL=double)
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "csrc", "go1sim.cu")
DEPS = [SRC, os.path.join(HERE, "csrc", "go1_warp.cuh"), os.path.join(HERE, "csrc", "go1_pair.cuh"), os.path.join(HERE, "csrc", "go1_policy_tc.cuh"),
        os.path.join(HERE, "csrc", "go1_render.cuh"), os.path.join(HERE, "csrc", "go1_model.h"),
        os.path.join(os.path.dirname(HERE), "include", "go1sim.h"), os.path.abspath(__file__)]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
              "-shared", "--expt-relaxed-constexpr"]


# fp32 product build: MUFU-based division / square root (<= 2 ulp) instead of the IEEE slow paths, denormals flushed to zero
# (drops the denormal fix-up code of those sequences: -2 % kernel time, parity unchanged; quantities here are never < 1e-38)
F32_FLAGS = ["-prec-div=false", "-prec-sqrt=false", "-ftz=true"]


def _stale(out: str) -> bool:
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(d) > t for d in DEPS)


def build(force: bool = False, verbose: bool = False, variants=("f32", "f64")) -> dict:
    nvcc = os.environ.get("NVCC", "nvcc")
    outs = {}
    for v in variants:
        out = os.path.join(HERE, "libgo1sim.so" if v == "f32" else "libgo1sim_f64.so")
        outs[v] = out
        if not force and not _stale(out):
            continue
        cmd = [nvcc] + NVCC_FLAGS + (["-DGO1_REAL=double", "-DGO1_MINBLOCKS=1"] if v == "f64" else F32_FLAGS) + \
              (["-Xptxas", "-v"] if verbose else []) + ["-o", out, SRC]
        print("[build]", " ".join(cmd), file=sys.stderr)
        subprocess.check_call(cmd)
    return outs


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
