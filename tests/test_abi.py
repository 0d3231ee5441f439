"""CPU-side checks of the drop-in boundary: libvnsgpu.so loads and exports exactly the entry
points include/vnsgpu.h declares (no compute calls - there is no GPU here)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "vnsgpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    text = re.sub(r"static inline[^{]*\{.*?\n\}", "", text, flags=re.S)     # header-only helpers are not exports
    return sorted(set(re.findall(r"\b(vnsgpu_[a-z_0-9]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    from vns_sop_b200 import engine
    lib = engine.load_library()
    syms = declared_symbols()
    assert len(syms) >= 30
    for s in syms:
        assert hasattr(lib, s), s
    assert sorted(engine.EXPORTS) == syms
    assert b"sm_100a" in lib.vnsgpu_version()


def test_rng_replay_matches_glibc(golden):
    """vnsgpu_rng_seed/next (host side of a16) against the reference process's srand/rand."""
    from vns_sop_b200 import engine
    for seed, vals in golden["_rand"].items():
        g = engine.seeded_rng(int(seed))
        assert [engine.rng_next(g) for _ in vals] == vals
    libc = ctypes.CDLL("libc.so.6")
    libc.srand(777)
    g = engine.seeded_rng(777)
    assert [engine.rng_next(g) for _ in range(5000)] == [libc.rand() for _ in range(5000)]


def test_no_gpu_fails_loudly():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from vns_sop_b200 import engine
    with pytest.raises(engine.VnsGpuError):
        engine.Engine(0)


def test_product_never_touches_the_oracle():
    pkg = os.path.join(ROOT, "vns-sop_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inl", ".h", ".cpp")):
                src = open(os.path.join(dirpath, f)).read()
                src = src.replace("oracle/vns_oracle.c:vo_dubins_length", "")
                assert "oracle" not in src, os.path.join(dirpath, f)


def test_header_is_plain_c_and_packers_match_python(tmp_path):
    """include/vnsgpu.h must compile as C (the boundary is a C ABI: no C++ or torch types), and its inline helpers
    (8-byte / 4-byte descriptors, result words) must agree with the Python mirror in engine.py bit for bit."""
    import subprocess
    import numpy as np
    from vns_sop_b200 import engine
    src = tmp_path / "t.c"
    src.write_text(r'''
#include <stdio.h>
#include "vnsgpu.h"
int main(void) {
   static const int c[][3] = {{0, 7, 0}, {1, 41, 29}, {3, 5, 17}, {4, 16383, 2}, {3, 0, 16383}, {0, 200, 1999}};
   for (unsigned k = 0; k < sizeof(c) / sizeof(c[0]); ++k) {
      const int gain = c[k][0] <= 1 ? c[k][1] : -1;
      printf("%llu %u\n", (unsigned long long)vnsgpu_pack_cand(c[k][0], c[k][1], c[k][2], gain),
             (unsigned)vnsgpu_pack_cand32(c[k][0], c[k][1], c[k][2], gain >= 0));
   }
   printf("%d %d %d %d\n", (int)vnsgpu_result_len(0x60002a37u), vnsgpu_result_flags(0x60002a37u),
          vnsgpu_result_flags(0xffffffffu), (int)sizeof(vnsgpu_cand));
   return 0;
}
''')
    exe = tmp_path / "t"
    subprocess.check_call(["gcc", "-std=c99", "-Wall", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    out = subprocess.check_output([str(exe)], text=True).split("\n")
    rows = np.array([[0, 7, 0], [1, 41, 29], [3, 5, 17], [4, 16383, 2], [3, 0, 16383], [0, 200, 1999]], np.int32)
    p64, p32 = engine.VnsSopPath.pack_cands(rows), engine.VnsSopPath.pack_cands32(rows)
    for k in range(len(rows)):
        a, b = out[k].split()
        assert int(a) == int(p64[k]) and int(b) == int(p32[k]), (k, out[k], p64[k], p32[k])
    assert out[len(rows)].split() == [str(0x2a37), "3", str(engine.F_BAD), "16"]


def test_rng_advance_equals_stepping():
    """vnsgpu_rng_advance (jump-ahead through powers of the companion matrix of x_i = x_{i-3} + x_{i-31}) against calling
    rand() that many times; host-only code, no GPU needed."""
    from vns_sop_b200 import engine
    for seed, n in ((1, 0), (1, 1), (7, 2), (12345, 31), (3, 512), (99, 100003)):
        a, b = engine.seeded_rng(seed), engine.seeded_rng(seed)
        for _ in range(n):
            engine.rng_next(a)
        engine.rng_advance(b, n)
        assert [engine.rng_next(a) for _ in range(40)] == [engine.rng_next(b) for _ in range(40)], (seed, n)
