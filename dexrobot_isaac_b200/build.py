"""Builds libdexstep.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

    python -m dexrobot_isaac_b200.build [--force] [--verbose]

nvcc cross-compiles without a GPU.  The .so is git-ignored but travels with the working tree.
Flags: -fmad=false and no fast-math on purpose (csrc/dexstep.cu header comment): thresholded
masks sit downstream of fp32 arithmetic that must round like the reference's separate torch ops.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
REPO_ROOT = os.path.dirname(PKG_DIR)
CSRC = os.path.join(PKG_DIR, "csrc")
SOURCES = [os.path.join(CSRC, "dexstep.cu")]
HEADERS = [os.path.join(REPO_ROOT, "include", "dexstep.h")] + sorted(
    os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cuh"))
LIB_PATH = os.path.join(PKG_DIR, "libdexstep.so")

NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-fmad=false", "-std=c++17",
              "-shared", "-Xcompiler", "-fPIC", "--cudart", "shared"]


def find_nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found (set NVCC=/path/to/nvcc)")


def needs_build() -> bool:
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    return any(os.path.getmtime(p) > t for p in SOURCES + HEADERS + [os.path.abspath(__file__)])


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not needs_build():
        return LIB_PATH
    cmd = [find_nvcc()] + NVCC_FLAGS + os.environ.get("DEXSTEP_NVCC_FLAGS", "").split() + ["-I", os.path.join(REPO_ROOT, "include")]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    cmd += SOURCES + ["-o", LIB_PATH]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed: " + " ".join(cmd))
    return LIB_PATH


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
