"""Builds libstepth_cuda.so in-tree with nvcc for sm_100a (cross-compiles without a GPU)."""
from __future__ import annotations

import glob
import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(_HERE, "libstepth_cuda.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-ffp-contract=off,-fno-fast-math",
    "-shared",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libstepth_cuda.so cannot be built (there is no CPU fallback)")


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")) + glob.glob(os.path.join(CSRC, "*.cuh")) + glob.glob(os.path.join(CSRC, "*.cpp"))) + [
        os.path.join(os.path.dirname(_HERE), "include", "stepth_cuda.h")
    ]


def is_stale() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.exists(s) and os.path.getmtime(s) > t for s in sources())


def build(force: bool = False, verbose: bool = False, out: str | None = None, defs=()) -> str:
    """`out` / `defs` build a tuning variant next to the product library (kernel sweeps; see _ffi.lib)."""
    if out is None and not force and not is_stale():
        return LIB_PATH
    out = out or LIB_PATH
    cmd = [_nvcc(), *NVCC_FLAGS, *[f"-D{d}" for d in defs], "-o", out, os.path.join(CSRC, "stepth_cuda.cu"),
           os.path.join(CSRC, "hostcopy.cpp")]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stdout + r.stderr)
    return out


DEBUG_LIB_PATH = os.path.join(_HERE, "libstepth_cuda_dbg.so")


def build_debug(force: bool = False) -> str:
    """the -DSTEPTH_DEBUG_CHECKS variant (index checks of the search kernels; tests/test_gpu_parity.py runs cases through it)"""
    if not force and os.path.exists(DEBUG_LIB_PATH) and all(
            not os.path.exists(s) or os.path.getmtime(s) <= os.path.getmtime(DEBUG_LIB_PATH) for s in sources()):
        return DEBUG_LIB_PATH
    return build(out=DEBUG_LIB_PATH, defs=["STEPTH_DEBUG_CHECKS=1"])


if __name__ == "__main__":
    print(build(force=True, verbose=True))
