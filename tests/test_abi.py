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
