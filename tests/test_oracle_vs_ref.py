"""Pins the C restatement to the reference's own code (oracle/_ref/liboracle_ref.so) on randomised
states beyond the committed vectors.  Runs where oracle/_ref was built (needs the instance as a
.sop text file: re-emitted from the fixture by instances.write_sop)."""
import numpy as np
import pytest

from conftest import SMALL, load_problem, random_candidates
from oracle import binding as ob
from vns_sop_b200 import instances as ins

pytestmark = pytest.mark.skipif(not ob.RefOracle.available(), reason="oracle/_ref not built")


@pytest.mark.parametrize("name", SMALL)
def test_random_walk_matches_reference(name, tmp_path):
    sf, pb = load_problem(name)
    f = str(tmp_path / (name + ".sop"))
    ins.write_sop(sf, f)
    r = ob.RefOracle(f)
    o = ob.Oracle(pb)
    assert (o.dense() == r.dense()).all()
    po, pr = o.path(), r.path()
    po.greedy()
    pr.greedy()
    rs = np.random.RandomState(5)
    seed = 100 + len(name)
    g = ob.new_rng(seed)
    r.seed(seed)
    for it in range(8):
        k = 1 + int(rs.randint(0, 2))
        po.shake(k, g)
        pr.shake(k)
        assert po.T == pr.T and (po.vns == pr.vns).all()
        cands = random_candidates(rs, po.vns, po.T, 200) if po.T >= 3 else None
        if cands is not None:
            a, b = po.eval_batch(cands), pr.eval_batch(cands)
            assert (a[0] == b[0]).all() and (a[1] == b[1]).all()
        n = int(rs.randint(50, 400))
        skip = int(rs.randint(0, 2))
        po.local_search(k, n, g, skip=skip)
        pr.local_search(k, n, skip=skip)
        assert po.T == pr.T and (po.vns == pr.vns).all() and po.length() == pr.length()
        Fo, Fr = po.layers(), pr.layers(o.mp)
        for l in range(po.T + 2):
            c = pb.start if l == 0 else (pb.goal if l == po.T + 1 else po.vns[l - 1])
            for x, y in zip(Fo, Fr):
                assert (x[l, :pb.m[c]] == y[l, :pb.m[c]]).all()
    assert ob.rng_next(g) == r.rand()


def test_real_valued_explicit_matrix_matches_reference(tmp_path):
    """Floating-point association pinned against the REFERENCE, not only against the port: a
    real-valued EXPLICIT matrix through the compiled reference and through the C port - table, greedy tour, 20 000
    candidates with their lower bounds, shake + local-search calls, generator position."""
    from conftest import real_valued_explicit_krob200
    sf, pb = real_valued_explicit_krob200()
    f = str(tmp_path / "krob200_real.sop")
    ins.write_sop(sf, f)
    r = ob.RefOracle(f)
    o = ob.Oracle(pb)
    W = o.dense()
    assert (W == r.dense()).all() and (W != np.round(W)).any()
    po, pr = o.path(), r.path()
    po.greedy()
    pr.greedy()
    assert po.T == pr.T and (po.vns == pr.vns).all() and po.length() == pr.length()
    rs = np.random.RandomState(3)
    cands = random_candidates(rs, po.vns, po.T, 20000)
    a, b = po.eval_batch(cands), pr.eval_batch(cands)
    assert (a[0] == b[0]).all() and (a[1] == b[1]).all()
    g = ob.new_rng(9)
    r.seed(9)
    for it in range(10):
        k = 1 + it % 2
        po.shake(k, g)
        pr.shake(k)
        po.local_search(k, 1600, g)
        pr.local_search(k, 1600)
        assert po.T == pr.T and (po.vns == pr.vns).all() and po.length() == pr.length() and po.reward() == pr.reward()
    assert ob.rng_next(g) == r.rand()
