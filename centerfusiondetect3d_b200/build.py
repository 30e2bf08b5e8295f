"""Builds libcfd3d_b200.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

    python -m centerfusiondetect3d_b200.build [--force] [--prof]

--prof builds libcfd3d_b200_prof.so with -DCFD3D_PROFILE (in-kernel phase clocks, tools/prof_phases.py; select it with
CFD3D_LIB=<path>); the product library carries no instrumentation.

nvcc cross-compiles without a GPU; the resulting .so travels to the GPU box with the repo snapshot.
Flags: -fmad=false keeps one rounding per written fp32/fp64 operation (parity contract),
-lineinfo maps ncu's source page back to the .cu files.
"""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libcfd3d_b200.so")
SOURCES = ["api.cu", "peak_ops.cu", "frustum.cu", "stage.cu", "post.cu"]
HEADERS = ["common.cuh", "select.cuh", "geometry.cuh", "kernels.cuh", "merge.cuh", "assoc.cuh", "paint.cuh", "post.cuh", os.path.join("..", "..", "include", "cfd3d_b200.h")]

NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-fmad=false",
    "-Xcompiler", "-fPIC,-fvisibility=hidden,-O2", "-shared", "-Xptxas", "-v",
]


LIB_PROF = os.path.join(HERE, "libcfd3d_b200_prof.so")


def _stale(lib: str = LIB) -> bool:
    if not os.path.isfile(lib):
        return True
    t = os.path.getmtime(lib)
    deps = [os.path.join(CSRC, s) for s in SOURCES + HEADERS] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force: bool = False, verbose: bool = False, prof: bool = False) -> str:
    lib = LIB_PROF if prof else LIB
    if not force and not _stale(lib):
        return lib
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-DCFD3D_PROFILE"] if prof else []) + ["-o", lib] + [os.path.join(CSRC, s) for s in SOURCES]
    res = subprocess.run(cmd, capture_output=True, text=True)
    log = os.path.join(HERE, "build_prof.log" if prof else "build.log")
    with open(log, "w") as f:
        f.write(" ".join(cmd) + "\n" + res.stdout + res.stderr)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed building libcfd3d_b200.so (see %s)" % log)
    if verbose:
        sys.stderr.write(res.stderr)
    return lib


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True, prof="--prof" in sys.argv))
