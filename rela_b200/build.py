"""Build the in-tree native libraries of rela_b200.

librela_b200.so   CUDA kernels + C ABI (nvcc, sm_100a only, no torch)
_rela_shim*.so    pybind module with rela's Python surface over the C ABI
                  (built by build_shim(); needs torch headers, slower)
"""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "librela_b200.so")
KERNEL_SOURCES = ["replay.cu", "row_copy.cu", "select.cu", "actor_math.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-Wall", "-shared", "-cudart", "static",
]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def nvcc_path():
    for cand in (os.environ.get("NVCC"), "/usr/local/cuda/bin/nvcc", "nvcc"):
        if cand and (os.path.isabs(cand) and os.path.exists(cand) or not os.path.isabs(cand)):
            return cand
    raise RuntimeError("nvcc not found")


def build_kernels(force=False, verbose=False):
    srcs = [os.path.join(CSRC, s) for s in KERNEL_SOURCES]
    deps = srcs + [os.path.join(CSRC, h) for h in ("kernels.h", "ptx.cuh")] + [
        os.path.join(ROOT, "include", "rela_b200.h")]
    if not force and not _newer(LIB, deps):
        return LIB
    cmd = [nvcc_path()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + srcs + ["-o", LIB]
    subprocess.run(cmd, check=True, cwd=CSRC)
    return LIB


if __name__ == "__main__":
    build_kernels(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
