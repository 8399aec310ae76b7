"""Builds libcrt.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

The flags are part of the numerics contract: no implicit fused multiply-add, IEEE
division and square root, denormals kept.
"""
from __future__ import annotations

import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libcrt.so")
SOURCES = ["crt_api.cu", "crt_kernels.cu", "crt_wavefront.cu", "crt_bvh_device.cu", "crt_bvh.cpp"]
HEADERS = ["crt_device.cuh", "crt_path.cuh", "crt_kernels.h", "crt_wavefront.h", "crt_bvh.h", "crt_bvh_ploc.h", "crt_bvh_device.h", "../../include/crt.h", "../../include/crt_detmath.h"]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC,-fvisibility=hidden,-O2,-ffp-contract=off,-pthread",
    "-shared", "-cudart", "static",
]


def nvcc_path() -> str:
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(os.path.join(CSRC, f)) > t for f in SOURCES + HEADERS)


def build_variant(out: str, defines: list[str]) -> str:
    """Tuning aid: the same library with extra -D defines, written to `out` (select it with CRT_LIBCRT=<path>)."""
    cmd = [nvcc_path(), *NVCC_FLAGS, *[f"-D{d}" for d in defines], "-o", out, *[os.path.join(CSRC, s) for s in SOURCES]]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + r.stdout + r.stderr)
    return out


def build_libcrt(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB
    cmd = [nvcc_path(), *NVCC_FLAGS, "-o", LIB, *[os.path.join(CSRC, s) for s in SOURCES]]
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stdout + r.stderr)
    return LIB


if __name__ == "__main__":
    import sys
    print(build_libcrt(force=True, verbose="-v" in sys.argv))
