import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

SMALL = ["berlin52", "krob200", "opn_s08", "dop_s04", "dop_s08", "dop_s12"]
ALL = SMALL + ["vm1084"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(GOLDEN_DIR, "golden.json")) as fh:
        return json.load(fh)


def load_problem(name):
    from vns_sop_b200 import instances as ins
    sf = ins.load_npz(os.path.join(GOLDEN_DIR, name + ".npz"))
    return sf, ins.build_problem(sf)


def load_arrays(name):
    return np.load(os.path.join(GOLDEN_DIR, name + "_golden.npz"))


def random_candidates(rs, vns, T, n):
    out = []
    outside = vns[T:]
    while len(out) < n:
        op = rs.randint(0, 5)
        if op == 0 and len(outside):
            out.append((0, int(rs.choice(outside)), rs.randint(0, T + 1)))
        elif op == 1 and len(outside) and T >= 1:
            out.append((1, int(rs.choice(outside)), rs.randint(0, T)))
        elif op == 2 and T >= 1:
            out.append((2, rs.randint(0, T), 0))
        elif op == 3 and T >= 2:
            a, b = rs.randint(0, T), rs.randint(0, T + 1)
            if abs(a - b) >= 2:
                out.append((3, a, b))
        elif op == 4 and T >= 3:
            a, b = rs.randint(0, T), rs.randint(0, T)
            if abs(a - b) >= 2:
                out.append((4, a, b))
    return np.array(out, dtype=np.int32)


def real_valued_explicit_krob200():
    """krob200 as a REAL-VALUED explicit matrix (raw Euclid + seeded U[0,1) noise, asymmetric): the one input on which the
    floating-point association of the reference's sums can show - the reference parses EXPLICIT
    weights as doubles (dataset_loader_sop.cpp:254-257), so the same matrix goes through the compiled reference, the C
    port and the GPU's IEEE double path."""
    import dataclasses
    from vns_sop_b200 import instances as ins
    sf, _ = load_problem("krob200")
    xy = np.asarray(sf.coords, np.float64)
    d = np.sqrt(((xy[:, None, :] - xy[None, :, :]) ** 2).sum(-1))
    W = d + np.random.RandomState(20190147).rand(*d.shape)
    np.fill_diagonal(W, 0.0)
    sf2 = dataclasses.replace(sf, coords=None, weights=W, edge_weight_type="EXPLICIT", name="krob200_real")
    return sf2, ins.build_problem(sf2)
