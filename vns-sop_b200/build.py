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
    "-Xcompiler", "-fPIC",
]
# translation units of the library (each compiled on its own core, then linked; no relocatable device code: no kernel
# calls a device function of another unit)
UNITS = ["vnsgpu.cu", "api_path.cu", "api_search.cu", "chain2.cu", "chain2_k8.cu", "chain2_k16.cu", "chain3.cu", "chain3_g.cu", "best.cu", "comm.cu", "noalloc.cu", "measure.cu", "gen.cu"]
OBJ = os.path.join(HERE, "build")


def sources():
    return [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))] + [os.path.join(HERE, "..", "include", "vnsgpu.h")]


def stale():
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    return any(os.path.getmtime(s) > t for s in sources())


def _unit_deps(unit):
    """headers a unit includes (transitively, by scanning #include "..." lines)"""
    seen, todo = set(), [unit]
    while todo:
        f = todo.pop()
        if f in seen:
            continue
        seen.add(f)
        path = os.path.normpath(os.path.join(CSRC, f))
        if not os.path.exists(path):
            continue
        for line in open(path):
            line = line.strip()
            if line.startswith('#include "'):
                todo.append(line.split('"')[1])
    return [os.path.normpath(os.path.join(CSRC, f)) for f in seen]


def build(force=False, verbose=False):
    # always decided per object (an edit made while an earlier build was still linking is older than the library)
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    os.makedirs(OBJ, exist_ok=True)
    units = [u for u in UNITS if os.path.exists(os.path.join(CSRC, u))]
    procs, objs = [], []
    for u in units:
        o = os.path.join(OBJ, u.replace(".cu", ".o"))
        objs.append(o)
        deps = _unit_deps(u)
        hdr = os.path.join(HERE, "..", "include", "vnsgpu.h")
        if not force and os.path.exists(o) and all(os.path.getmtime(d) <= os.path.getmtime(o) for d in deps + [hdr] if os.path.exists(d)):
            continue
        cmd = [nvcc] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", o, os.path.join(CSRC, u)]
        log = open(o + ".log", "w")
        procs.append((u, subprocess.Popen(cmd, cwd=CSRC, stdout=log, stderr=subprocess.STDOUT), log))
    failed = []
    for u, p, log in procs:
        rc = p.wait()
        log.close()
        if rc != 0:
            failed.append(u)
    if verbose or failed:
        for u, p, log in procs:
            print("---- %s ----" % u)
            print(open(log.name).read())
    if failed:
        raise RuntimeError("nvcc failed on " + ", ".join(failed))
    if procs or not os.path.exists(LIB) or any(os.path.getmtime(o) > os.path.getmtime(LIB) for o in objs):
        subprocess.check_call([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs, cwd=CSRC)
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
           os.path.join(HOST, "sop_vns_b200.cpp"), "-o", CLI, "-L", HERE, "-lvnsgpu", "-pthread", "-Wl,-rpath,$ORIGIN"]
    subprocess.check_call(cmd)
    return CLI


if __name__ == "__main__":
    print(build(force=True, verbose=True))
    print(build_cli(force=True))
