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
KERNEL_SOURCES = ["replay.cu", "row_copy.cu", "gather_convert.cu", "select.cu", "actor_math.cu", "group.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC,-Wall", "-shared", "-cudart", "static", "-lrt",
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


def _torch_flags():
    import sysconfig

    import torch

    tdir = os.path.dirname(torch.__file__)
    inc = [os.path.join(tdir, "include"), os.path.join(tdir, "include", "torch", "csrc", "api", "include"),
           sysconfig.get_paths()["include"], "/usr/local/cuda/include"]
    cxx = ["-std=c++17", "-O2", "-fPIC", "-Wall", "-Wno-unused-parameter", "-Wno-sign-compare",
           "-Wno-deprecated-declarations", "-Wno-attributes",
           "-D_GLIBCXX_USE_CXX11_ABI=%d" % int(torch._C._GLIBCXX_USE_CXX11_ABI)]
    cxx += ["-I" + i for i in inc]
    libdir = os.path.join(tdir, "lib")
    ld = ["-L" + libdir, "-Wl,-rpath," + libdir, "-ltorch", "-ltorch_cpu", "-lc10", "-lc10_cuda", "-ltorch_cuda",
          "-ltorch_python"]
    return cxx, ld


CPP = os.path.join(HERE, "cpp")
SHIM = os.path.join(HERE, "_C.so")
SHIM_SOURCES = ["replay_core.cc", "sharded_core.cc", "issuer_core.cc", "sequence_assembler.cc", "pybind_module.cc"]


def build_shim(force=False):
    """rela_b200/_C.so: the C++ host layer (cpp/) + the pybind module with rela's
    replay classes, linked against librela_b200.so (found through $ORIGIN)."""
    build_kernels()
    srcs = [os.path.join(CPP, s) for s in SHIM_SOURCES]
    hdrs = [os.path.join(CPP, "rela_b200", h) for h in os.listdir(os.path.join(CPP, "rela_b200"))]
    deps = srcs + hdrs + [os.path.join(ROOT, "include", "rela_b200.h")]
    if not force and not _newer(SHIM, deps):
        return SHIM
    cxx, ld = _torch_flags()
    cxx += ["-I" + CPP, "-I" + os.path.join(ROOT, "include")]
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    procs, objs = [], []
    for s in srcs:
        o = os.path.join(objdir, os.path.basename(s) + ".o")
        objs.append(o)
        if force or _newer(o, [s] + hdrs + [deps[-1]]):
            procs.append(subprocess.Popen(["g++"] + cxx + ["-c", s, "-o", o]))
    for p in procs:
        if p.wait() != 0:
            raise RuntimeError("compiling the rela_b200 host layer failed")
    subprocess.run(["g++", "-shared", "-o", SHIM] + objs + ["-L" + HERE, "-lrela_b200", "-Wl,-rpath,$ORIGIN"] + ld,
                   check=True)
    return SHIM


if __name__ == "__main__":
    build_kernels(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
    if "--shim" in sys.argv:
        print(build_shim(force="--force" in sys.argv))
