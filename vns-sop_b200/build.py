"""In-tree build of libvnsgpu.so (CUDA kernels + C ABI) for sm_100a.  nvcc cross-compiles without a
GPU; the .so is git-ignored but travels to the GPU box with the repo snapshot."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libvnsgpu.so")

NVCC_FLAGS = [
    "-O3", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo",
    "-fmad=false",          # parity: no FMA contraction (dx*dx+dy*dy in the table build, Dubins terms)
    "-Xcompiler", "-fPIC", "-shared",
]


def sources():
    return [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))] + [os.path.join(HERE, "..", "include", "vnsgpu.h")]


def stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(s) > t for s in sources())


def build(force=False, verbose=False):
    if not force and not stale():
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    # VNSGPU_SPLIT_COMPILE=1: NVVM + ptxas of this one translation unit on all host cores (80 s instead of 4 min) - for
    # development only: the split changes register allocation (the latency-bound chain kernels lose ~10 %)
    split = ["-split-compile", "0"] if os.environ.get("VNSGPU_SPLIT_COMPILE") == "1" else []
    cmd = [nvcc] + NVCC_FLAGS + split + (["-Xptxas", "-v"] if verbose else []) + ["-o", LIB, os.path.join(CSRC, "vnsgpu.cu")]
    subprocess.check_call(cmd, cwd=CSRC)
    return LIB


HOST = os.path.join(HERE, "host")
CLI = os.path.join(HERE, "sop_vns_b200")


def build_cli(force=False):
    """The reference-compatible command line (host/sop_vns_b200.cpp) linked against libvnsgpu.so."""
    srcs = [os.path.join(HOST, f) for f in sorted(os.listdir(HOST))] + [os.path.join(HERE, "..", "include", f) for f in ("vnsgpu.h", "vnsgpu_path.hpp")]
    if not force and os.path.exists(CLI) and all(os.path.getmtime(s) <= os.path.getmtime(CLI) for s in srcs + [LIB]):
        return CLI
    cxx = os.environ.get("VNSGPU_CXX", "/usr/bin/g++")
    cmd = [cxx, "-O2", "-std=c++17", "-Wall", "-I", os.path.join(HERE, "..", "include"), "-I", HOST,
           os.path.join(HOST, "sop_vns_b200.cpp"), "-o", CLI, "-L", HERE, "-lvnsgpu", "-Wl,-rpath,$ORIGIN"]
    subprocess.check_call(cmd)
    return CLI


if __name__ == "__main__":
    print(build(force=True, verbose=True))
    print(build_cli(force=True))
