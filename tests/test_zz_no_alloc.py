"""vnsgpu_eval_batch_noalloc (scratch in a static __device__ array, no allocation) against vnsgpu_eval_batch.
Collected last on purpose; passes alone in a fresh process and after the other tests."""
import numpy as np
import pytest

from conftest import load_problem, random_candidates


@pytest.mark.gpu
def test_noalloc_eval_matches_eval_batch_twice():
    from vns_sop_b200 import engine
    sf, pb = load_problem("krob200")
    e = engine.Engine(0).load_problem(pb)
    try:
        rs = np.random.RandomState(5)
        p = e.path()
        p.greedy_insertion()
        T, vns = p.getNumTargets(), np.array(p.vns)
        # two calls in a row, different batches and sizes: the second finds the first call's contents in the array
        for n in (3000, 1234):
            cands = random_candidates(rs, vns, T, n)
            cands[::97, 2] = 9999                                     # out-of-range indices: NaN + F_BAD on both paths
            ln0, lb0, fl0, first0 = p.eval_batch(cands)
            ln1, lb1, fl1, first1 = p.eval_batch_noalloc(cands)
            ok = fl0 != engine.F_BAD
            assert (fl0 == fl1).all() and first0 == first1
            # bit-exact (the tolerance every batched-evaluation test here uses): same arithmetic in the reference's association
            assert (ln0[ok] == ln1[ok]).all() and (lb0[ok] == lb1[ok]).all()
            assert np.isnan(ln1[~ok]).all() and (~ok).sum() > 0
            # the search moves on between the calls
            rng = engine.seeded_rng(11)
            p.shake(1, rng)
            T, vns = p.getNumTargets(), np.array(p.vns)
        # beyond the fixed maximum: a clear error, nothing evaluated
        big = random_candidates(rs, vns, T, 16385)
        with pytest.raises(engine.VnsGpuError) as ei:
            p.eval_batch_noalloc(big)
        assert "VNSGPU_NOALLOC_MAX_CANDS" in str(ei.value)
        assert e.lib.vnsgpu_noalloc_max_cands() == 16384
        ln0, lb0, fl0, first0 = p.eval_batch(big[:16384])
        ln1, lb1, fl1, first1 = p.eval_batch_noalloc(big[:16384])       # exactly the maximum
        assert (ln0 == ln1).all() and (lb0 == lb1).all() and (fl0 == fl1).all() and first0 == first1
    finally:
        e.close()
