"""Build libadapt_b200.so (the C-ABI stage executor) for sm_100a with nvcc, in-tree.

Run as a script (`python csrc/build.py`) or through `__graft_entry__.build()`.  The shared
object sits next to this file so that it travels with a gpurun snapshot; it is git-ignored.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "libadapt_b200.so")
TRACE_LIB = os.path.join(HERE, "libadapt_b200_trace.so")  # developer build with the role timeline compiled in
SOURCES = ["adapt_b200.cu"]
HEADERS = ["conv_gemm_sm100.cuh", "conv_chain_sm100.cuh", "stem_pool_sm100.cuh", "ptx_sm100.cuh", "simt_kernels.cuh", os.path.join("..", "..", "include", "adapt_b200.h")]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; cannot build libadapt_b200.so")


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    for f in SOURCES + HEADERS:
        if os.path.getmtime(os.path.join(HERE, f)) > t:
            return True
    return False


def build(force: bool = False, verbose: bool = False) -> str:
    trace = os.environ.get("ADAPT_BUILD_TRACE", "0") == "1"
    if not trace and not force and not needs_build():
        return LIB
    out = TRACE_LIB if trace else LIB
    cmd = [
        _nvcc(),
        "-gencode", "arch=compute_100a,code=sm_100a",
        "-lineinfo", "-O3", "-std=c++17",
        "--shared", "-Xcompiler", "-fPIC",
        "-o", out,
    ] + (["-DADAPT_ENABLE_TRACE=1"] if trace else []) + os.environ.get("ADAPT_NVCC_FLAGS", "").split() + SOURCES
    if verbose:
        cmd.insert(1, "-Xptxas")
        cmd.insert(2, "-v")
    proc = subprocess.run(cmd, cwd=HERE, capture_output=True, text=True)
    if proc.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + proc.stdout + proc.stderr)
    if verbose:
        sys.stderr.write(proc.stderr)
    return out


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
