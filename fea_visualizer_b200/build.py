"""Builds libfeab.so (the CUDA kernels + C ABI) in-tree with nvcc for sm_100a.  No torch involved."""
from __future__ import annotations

import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libfeab.so")
SOURCES = ["feab.cu", "surface.cu", "noalloc.cu", "ply_writer.cpp"]


def _deps():
    """Everything libfeab.so is built from: every file under csrc/ and the C-ABI headers (globbed, so that a new
    header can never be forgotten here)."""
    inc = os.path.join(HERE, "..", "include")
    out = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith((".cu", ".cuh", ".cpp", ".hpp", ".h"))]
    out += [os.path.join(inc, f) for f in sorted(os.listdir(inc)) if f.endswith(".h")]
    return out


NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC,-ffp-contract=off", "-shared"]


HASHFILE = LIB + ".srchash"


def source_hash() -> str:
    """sha256 over the contents of every build input + the flags: libfeab.so is rebuilt whenever this differs from the
    hash recorded next to it (mtimes do not survive a checkout or the copy to the GPU box)."""
    import hashlib
    h = hashlib.sha256(" ".join(NVCC_FLAGS).encode())
    for f in _deps():
        h.update(os.path.basename(f).encode())
        h.update(open(f, "rb").read())
    return h.hexdigest()


def stale() -> bool:
    if not os.path.exists(LIB) or not os.path.exists(HASHFILE):
        return True
    return open(HASHFILE).read().strip() != source_hash()


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not stale():
        return LIB
    nvcc = os.environ.get("NVCC", "nvcc")
    cmd = [nvcc, *NVCC_FLAGS, "-o", LIB, *[os.path.join(CSRC, s) for s in SOURCES]]
    if verbose:
        cmd += ["-Xptxas", "-v"]
        print(" ".join(cmd))
    out = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or out.returncode != 0:
        sys.stderr.write(out.stdout + out.stderr)
    if out.returncode != 0:
        raise RuntimeError("nvcc failed building libfeab.so")
    with open(HASHFILE, "w") as f:
        f.write(source_hash() + "\n")
    return LIB


ROOT = os.path.dirname(HERE)
CPP_DRIVER_SRC = os.path.join(ROOT, "tests", "cpp", "cuda_solver_driver.cpp")
CPP_DRIVER = os.path.join(ROOT, "tests", "cpp", "cuda_solver_driver")


def build_cpp_host(force: bool = False) -> str:
    """g++ build of the headless C++ host program that drives include/FEM/CUDASolver.hpp (the C++ mirror of the
    reference's Solver interface) against libfeab.so.  Host C++ only; links the C ABI, nothing else."""
    inc = os.path.join(ROOT, "include")
    deps = [CPP_DRIVER_SRC, LIB] + [os.path.join(inc, "FEM", f) for f in os.listdir(os.path.join(inc, "FEM"))] + \
           [os.path.join(inc, "feab.h"), os.path.join(inc, "Utils", "Surface.hpp")]
    if not force and os.path.exists(CPP_DRIVER) and all(os.path.getmtime(d) <= os.path.getmtime(CPP_DRIVER) for d in deps):
        return CPP_DRIVER
    cmd = [os.environ.get("CXX", "g++"), "-std=c++17", "-O2", "-Wall", "-I", inc, CPP_DRIVER_SRC, "-o", CPP_DRIVER, "-L", HERE, "-lfeab",
           "-Wl,-rpath,$ORIGIN/../../fea_visualizer_b200"]
    out = subprocess.run(cmd, capture_output=True, text=True)
    if out.returncode != 0:
        sys.stderr.write(out.stdout + out.stderr)
        raise RuntimeError("g++ failed building tests/cpp/cuda_solver_driver")
    return CPP_DRIVER


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose=True)
    print(LIB)
    print(build_cpp_host(force="--force" in sys.argv))
