#!/usr/bin/env python
"""Build the sm_100a kernel library in-tree: csrc/*.cu -> p2pfl_b200/lib/libp2pfl_b200.so.

nvcc cross-compiles for sm_100a without a GPU; the built .so is git-ignored but
travels to the GPU box with the repo snapshot.
"""

from __future__ import annotations

import glob
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(ROOT, "p2pfl_b200", "lib", "libp2pfl_b200.so")
ORACLE_SRC = os.path.join(ROOT, "csrc", "oracle", "oracle.cpp")
ORACLE_OUT = os.path.join(ROOT, "p2pfl_b200", "lib", "libp2pfl_oracle.so")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]


def sources() -> list:
    return sorted(glob.glob(os.path.join(ROOT, "csrc", "*.cu")))


def needs_build() -> bool:
    if not os.path.exists(OUT):
        return True
    t = os.path.getmtime(OUT)
    deps = sources() + glob.glob(os.path.join(ROOT, "csrc", "*.cuh")) + glob.glob(os.path.join(ROOT, "include", "*.h"))
    return any(os.path.getmtime(d) > t for d in deps)


def build_oracle(force: bool = False) -> str:
    """The plain C++ oracles / timed CPU baselines (csrc/oracle): g++, no CUDA."""
    if not force and os.path.exists(ORACLE_OUT) and os.path.getmtime(ORACLE_OUT) >= os.path.getmtime(ORACLE_SRC):
        return ORACLE_OUT
    os.makedirs(os.path.dirname(ORACLE_OUT), exist_ok=True)
    res = subprocess.run([shutil.which("g++") or "g++", "-O2", "-std=c++17", "-fPIC", "-shared", ORACLE_SRC, "-o", ORACLE_OUT],
                         capture_output=True, text=True)  # fmt: skip
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("g++ failed on the oracle library")
    return ORACLE_OUT


def build(force: bool = False, verbose: bool = False, bringup: bool = False) -> str:
    """``bringup``: compile the per-link clock stamps and the mbarrier-wait trap into the GEMM kernels
    (-DP2PFL_BRINGUP, for experiments with tools/profile_target.py only; never for a measured run)."""
    build_oracle(force)
    if not force and not needs_build():
        return OUT
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    os.makedirs(os.path.dirname(OUT), exist_ok=True)
    cmd = [nvcc, *ARCH, "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC", "-shared"]
    if verbose:
        cmd += ["-Xptxas", "-v"]
    if bringup:
        cmd += ["-DP2PFL_BRINGUP"]
    cmd += sources() + ["-o", OUT]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("nvcc failed")
    if verbose:
        sys.stderr.write(res.stderr)
    return OUT


if __name__ == "__main__":
    path = build(force="--force" in sys.argv or "--bringup" in sys.argv, verbose="-v" in sys.argv, bringup="--bringup" in sys.argv)
    print(path)
