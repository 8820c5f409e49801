"""Builds libsmcphys.so (CUDA, sm_100a only) in-tree with nvcc.  No CPU build exists."""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "libsmcphys.so")
SOURCES = ["smcphys.cu"]
DEPS = ["smcphys.cu", "physics.cuh", "rng.cuh", "nccl_dl.h", os.path.join("..", "..", "include", "smcphys.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    # approximate sqrt / division (MUFU) and flush-to-zero in the float32 physics; sin/cos/log stay accurate
    "-prec-sqrt=false", "-prec-div=false", "-ftz=true",
    "--shared", "-Xcompiler", "-fPIC", "-ldl",
]


def nvcc_path():
    p = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(p):
        raise RuntimeError("nvcc not found: libsmcphys can only be built with the CUDA toolkit")
    return p


def stale():
    if not os.path.exists(SO):
        return True
    t = os.path.getmtime(SO)
    return any(os.path.getmtime(os.path.join(CSRC, d)) > t for d in DEPS) or os.path.getmtime(__file__) > t


def build(force=False, verbose=False):
    if not force and not stale():
        return SO
    cmd = [nvcc_path()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-o", SO] + SOURCES
    subprocess.check_call(cmd, cwd=CSRC)
    return SO


if __name__ == "__main__":
    import sys
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
