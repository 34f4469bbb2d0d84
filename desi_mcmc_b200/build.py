"""Builds libceleste_b200.so (sm_100a) in-tree with nvcc.

The shared library lives next to this file so that it travels with the
repository snapshot to the GPU box; it is git-ignored.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG)
SRC = os.path.join(PKG, "csrc", "celeste_b200.cu")
HDR = os.path.join(ROOT, "include", "celeste_b200.h")
LIB = os.path.join(PKG, "libceleste_b200.so")
LIB_CHECKED = os.path.join(PKG, "libceleste_b200_checked.so")   # -DCG_CHECKED: bounds asserts in the kernels

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    # every FMA in the kernels is explicit; the float64 preparation must follow
    # the reference's unfused NumPy arithmetic
    "-fmad=false",
    "-shared", "-Xcompiler", "-fPIC",
]


def _nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; cannot build libceleste_b200.so")


def needs_build() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    csrc = os.path.dirname(SRC)
    deps = [HDR] + [os.path.join(csrc, f) for f in os.listdir(csrc)]
    return any(os.path.getmtime(p) > t for p in deps)


def build(force: bool = False, verbose: bool = False, ptxas_info: bool = False, checked: bool = False) -> str:
    if checked:
        cmd = [_nvcc(), *NVCC_FLAGS, "-DCG_CHECKED", "-I", os.path.join(ROOT, "include"), SRC, "-o", LIB_CHECKED]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
        return LIB_CHECKED
    if not force and not needs_build():
        return LIB
    cmd = [_nvcc(), *NVCC_FLAGS, "-I", os.path.join(ROOT, "include"), SRC, "-o", LIB]
    if ptxas_info:
        cmd[1:1] = ["-Xptxas", "-v"]
    if verbose:
        print(" ".join(cmd))
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    if verbose or ptxas_info:
        sys.stderr.write(res.stderr)
    return LIB


if __name__ == "__main__":
    if "--checked" in sys.argv:
        print(build(checked=True))
    else:
        print(build(force=True, verbose=True, ptxas_info="-v" in sys.argv))
