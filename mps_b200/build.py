"""Builds libmps_b200.so (the CUDA library behind include/mps_b200.h) IN-TREE with nvcc for
sm_100a, and the host C++ programs. Used by __graft_entry__.build(); also runnable:
    python -m mps_b200.build
"""
from __future__ import annotations

import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB = os.path.join(HERE, "libmps_b200.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-fmad=false",            # parity: every fp64 op individually rounded
    "-Xcompiler", "-fPIC", "-shared",
    "-ldl",                   # NCCL is bound with dlopen (mpsb_comm.cu)
]


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build_library(force=False, verbose=False, out=None, extra=()):
    """out / extra: experiment builds (tools/): another file name and extra -D flags."""
    if out is not None:
        srcs = sorted(glob.glob(os.path.join(HERE, "csrc", "*.cu")))
        nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
        subprocess.run([nvcc] + NVCC_FLAGS + list(extra) + ["-o", out] + srcs, check=True)
        return out
    srcs = sorted(glob.glob(os.path.join(HERE, "csrc", "*.cu")))
    deps = srcs + glob.glob(os.path.join(HERE, "csrc", "*.cuh")) + [os.path.join(ROOT, "include", "mps_b200.h")]
    if not force and not _newer(LIB, deps):
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB] + srcs
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("nvcc failed building libmps_b200.so")
    if verbose:
        print(r.stderr)
    return LIB


if __name__ == "__main__":
    print(build_library(force=True, verbose="-v" in sys.argv))


def build_host(force=False):
    """Host C++ programs (mps_b200/host/Makefile), if present."""
    mk = os.path.join(HERE, "host", "Makefile")
    if not os.path.exists(mk):
        return None
    r = subprocess.run(["make", "-C", os.path.join(HERE, "host")] + (["-B"] if force else []),
                       capture_output=True, text=True)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("host build failed")
    return os.path.join(HERE, "host")
