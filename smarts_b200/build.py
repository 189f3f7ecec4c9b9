"""Build ``libsmartsb200.so`` in-tree with nvcc for sm_100a (no torch JIT, no cmake).

``-fmad=false``: the numeric contract of include/sb_detmath.h forbids FMA contraction so that
results are bit-identical to the CPU oracle.  ``-lineinfo`` keeps ncu source pages usable.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libsmartsb200.so")
SOURCES = ["sb_api.cu"]
HOST_SOURCES = ["sb_hostxfer.cpp"]  # g++ (AVX-512 target attributes), linked into the same library
DEPS = ["sb_api.cu", "sb_hostxfer.cpp", "sb_events.cuh", "sb_roadwp.cuh", "sb_ttc.cuh", "sb_extras.cuh", "sb_lidar.cuh", "sb_kernels.cuh", "sb_control.cuh", "sb_waypoints.cuh", "sb_device.cuh", "sb_raster.cuh",
        os.path.join("..", "..", "include", "smarts_b200.h"), os.path.join("..", "..", "include", "sb_detmath.h"), os.path.join("..", "..", "include", "sb_lidar_table.h"), os.path.join("..", "..", "include", "sb_vehicle.h")]


def nvcc_path() -> str:
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", shutil.which("nvcc")):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found")


def is_stale() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    return any(os.path.getmtime(os.path.join(CSRC, d)) > t for d in DEPS)


def build(force: bool = False, verbose: bool = False, out: str = None, defines=()) -> str:
    """``out`` / ``defines`` build a variant next to the product library (kernel A/B experiments under profiles/;
    loaded with SMARTS_B200_LIB=<path>)."""
    if out is None and not force and not is_stale():
        return out or OUT
    objs = []
    for src in HOST_SOURCES:
        obj = os.path.join(CSRC, os.path.splitext(src)[0] + ".o")
        res = subprocess.run(["g++", "-O3", "-std=c++17", "-fPIC", "-pthread", "-c", os.path.join(CSRC, src), "-o", obj],
                             capture_output=True, text=True)
        if res.returncode != 0:
            sys.stderr.write(res.stdout + res.stderr)
            raise RuntimeError("g++ failed")
        objs.append(obj)
    cmd = [
        nvcc_path(), "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
        "-fmad=false", "--expt-extended-lambda", "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-shared",
        "-Xptxas", "-v" if verbose else "-O3",
        "-o", out or OUT,
    ] + ["-D" + d for d in defines] + [os.path.join(CSRC, s) for s in SOURCES] + objs
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed")
    if verbose:
        sys.stderr.write(res.stderr)
    return out or OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
