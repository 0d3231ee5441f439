"""Synthetic scaled instances (BASELINE.json config 5; vns-sop_b200/synthetic.py).  CPU: the generator is
deterministic, produces a well-formed .sop and survives a write/parse round trip.  GPU: a 2000-node /
200-cluster instance (same construction as the 20k-node bench workload, small enough for the CPU oracle)
against the oracle - table, bounds, DP vectors, every tryTo* / lower bound on long tours, a seeded
shake + local-search sequence - and size-independent properties of the 20k-node instance itself."""
import numpy as np
import pytest

from conftest import random_candidates


def test_generator_is_deterministic_and_well_formed(tmp_path):
    from vns_sop_b200 import instances as ins, synthetic as sy
    a, b = sy.generate(2000, 10), sy.generate(2000, 10)
    assert (a.coords == b.coords).all() and a.tmax == b.tmax
    assert all((x == y).all() for x, y in zip(a.set_nodes, b.set_nodes))
    assert len(a.set_nodes) == 201 and len(a.set_nodes[0]) == 1 and a.start_set == a.end_set == 0
    allnodes = np.concatenate(a.set_nodes)
    assert sorted(allnodes.tolist()) == list(range(2001))          # every node in exactly one set
    assert (a.set_profit[1:] == 10).all() and a.set_profit[0] == 0  # rule p1: profit = cluster size
    f = str(tmp_path / "s.sop")
    ins.write_sop(a, f)
    c = ins.parse_sop(f)
    assert (c.coords == a.coords).all() and c.tmax == a.tmax and c.edge_weight_type == "CEIL_2D"
    pb = ins.build_problem(c)
    assert pb.C == 202 and pb.N == 2002 and pb.m.max() == 10
    vns, T = sy.centroid_tour_state(pb)
    assert sorted(vns.tolist()) == list(range(1, 201)) and 20 < T < 200
    r = sy.generate(2000, 10, random_clusters=True)
    assert r.name.endswith("_RND") and not all((x == y).all() for x, y in zip(a.set_nodes, r.set_nodes))
    # proximity clusters are compact, random ones are not
    spread = lambda sf: np.mean([sf.coords[s].std(0).sum() for s in sf.set_nodes[1:]])
    assert spread(a) < 0.25 * spread(r)


@pytest.fixture(scope="module")
def synth():
    from oracle import binding as ob
    from vns_sop_b200 import engine, instances as ins, synthetic as sy
    sf = sy.generate(2000, 10)
    pb = ins.build_problem(sf)
    e = engine.Engine(0).load_problem(pb)
    o = ob.Oracle(pb)
    yield e, o, pb, sy.centroid_tour_state(pb)
    e.close()


@pytest.mark.gpu
def test_synthetic_table_bounds_and_dp(synth):
    e, o, pb, (vns, T) = synth
    assert (e.table_dense() == o.dense()).all()
    lo, hi = e.bounds()
    olo, ohi = o.bounds()
    assert (lo == olo).all() and (hi == ohi).all()
    pg, po = e.path(vns, T), o.path(vns, T)
    assert pg.getPathLength() == po.length() and pg.getPathLengthRear() == po.length_rear() and pg.getReward() == po.reward()
    F, B = pg.layers()
    oF, oB = po.layers()[:2]
    for l in range(T + 2):
        k = 1 if l in (0, T + 1) else 10
        assert (F[l, :k] == oF[l, :k]).all() and (B[l, :k] == oB[l, :k]).all(), l


@pytest.mark.gpu
@pytest.mark.parametrize("int_path", [True, False])
def test_synthetic_long_tour_candidates(synth, int_path):
    """T = O(100) targets: MOVE / EXCHANGE chains of up to ~T blocks, with and without the segment products."""
    e, o, pb, (vns, T) = synth
    pg, po = e.path(vns, T), o.path(vns, T)
    rs = np.random.RandomState(5)
    cands = random_candidates(rs, vns, T, 170000)              # >= C^2 / 4: the endpoint-vector form engages
    oln, olb = po.eval_batch(cands)
    e.set_int_path(int_path)
    try:
        # segment products: 1 = default build, 2 = one-level chains, 3 = two-level (tiles of 32 layers) whatever the size
        for seg, vec in ((1, True), (3, True), (1, 3), (3, 3), (2, False), (3, False), (1, False), (0, False)):      # vec 3: unpacked vector build, 16 lanes per product in the cross products
            e.set_segment_cache(seg)
            e.set_vector_cache(vec)
            ln, lb, fl, first = pg.eval_batch(pg.make_cands(cands))
            assert (ln == oln).all() and (lb == olb).all(), (int_path, seg, vec)
    finally:
        e.set_int_path(True)
        e.set_segment_cache(True)
        e.set_vector_cache(True)


@pytest.mark.gpu
def test_synthetic_shake_and_local_search(synth):
    from oracle import binding as ob
    from vns_sop_b200 import engine
    e, o, pb, (vns, T) = synth
    pg, po = e.path(vns, T), o.path(vns, T)
    pg.fitDOPtoBudget()
    po.fit()
    rg, ro = engine.seeded_rng(11), ob.new_rng(11)
    for k in (1, 2, 1, 2):
        pg.shake(k, rg)
        po.shake(k, ro)
        assert (pg.vns == po.vns).all() and pg.getNumTargets() == po.T
        pg._ls(k, 4000, rg, None, 1)
        po.local_search(k, 4000, ro)
        assert (pg.vns == po.vns).all() and pg.getNumTargets() == po.T
        assert pg.getPathLength() == po.length() and pg.getReward() == po.reward()
    assert engine.rng_next(rg) == ob.rng_next(ro)


@pytest.mark.gpu
def test_synthetic_600_clusters_large_instance_builds_vs_oracle():
    """6000 nodes / 600 clusters, T = 237: the smallest instance on which the per-state builds take their large-instance
    branches (from 512 clusters on: int32 row copies, lower-bound sums in k_vec_lbs; more than 256 layers' worth of clusters:
    two-level segment build with the 10-lane cross products) and the oracle still answers in seconds."""
    from oracle import binding as ob
    from vns_sop_b200 import engine, instances as ins, synthetic as sy
    pb = ins.build_problem(sy.generate(6000, 10))
    assert pb.C >= 512
    e = engine.Engine(0).load_problem(pb)
    try:
        vns, T = sy.centroid_tour_state(pb)
        po, pg = ob.Oracle(pb).path(vns, T), e.path(vns, T)
        assert pg.getPathLength() == po.length() and pg.getReward() == po.reward()
        cands = random_candidates(np.random.RandomState(7), vns, T, 120000)      # 4 n >= C^2: the endpoint-vector form engages
        oln, olb = po.eval_batch(cands)
        for seg, vec in ((1, True), (1, 3), (2, True), (3, 3), (1, False)):
            e.set_segment_cache(seg)
            e.set_vector_cache(vec)
            ln, lb, fl, first = pg.eval_batch(pg.make_cands(cands))
            assert (ln == oln).all() and (lb == olb).all(), (seg, vec)
        # a state change between two batches: the builds follow the path
        e.set_segment_cache(True)
        e.set_vector_cache(True)
        rg, ro = engine.seeded_rng(3), ob.new_rng(3)
        pg.shake(1, rg)
        po.shake(1, ro)
        assert (pg.vns == po.vns).all() and pg.getNumTargets() == po.T
        cands = random_candidates(np.random.RandomState(8), po.vns, po.T, 120000)
        oln, olb = po.eval_batch(cands)
        ln, lb, fl, first = pg.eval_batch(pg.make_cands(cands))
        assert (ln == oln).all() and (lb == olb).all()
        # greedy_insertion as batched rounds: from 512 clusters on its tryToAdd-only batches take the endpoint-vector form
        # without pair vectors; same tour as the one-block greedy (the reference's loop), with and without the vectors
        p1, p2, p3 = e.path(), e.path(), e.path()
        p1.greedy_insertion()
        rounds, evals = p2.greedy_insertion_batched()
        assert rounds == p1.getNumTargets() == p2.getNumTargets() and (p1.vns == p2.vns).all()
        assert p1.getPathLength() == p2.getPathLength() and p1.getReward() == p2.getReward()
        e.set_vector_cache(0)
        p3.greedy_insertion_batched()
        e.set_vector_cache(True)
        assert (p3.vns == p1.vns).all() and p3.getPathLength() == p1.getPathLength()
    finally:
        e.close()


@pytest.mark.gpu
def test_20k_instance_properties():
    """The bench workload itself (3.2 GB table, HBM regime): no oracle at this size inside the test budget, so
    size-independent properties - the table's block-major gather equals direct evaluation on sampled entries,
    forward == backward length, REMOVE then ADD-back identities, lower bound <= value, exact-int and double
    evaluation agree bit for bit, segment products on/off agree."""
    from vns_sop_b200 import engine, instances as ins, synthetic as sy
    sf = sy.generate()
    pb = ins.build_problem(sf)
    e = engine.Engine(0).load_problem(pb)
    try:
        vns, T = sy.centroid_tour_state(pb)
        p = e.path(vns, T)
        assert p.getPathLength() == p.getPathLengthRear() > 0
        cl, nd, calc = p.getPath()
        assert calc == p.getPathLength()                      # back-pointer walk re-sums to the DP value
        xy = pb.xy[pb.off[cl] + nd]
        direct = np.ceil(np.sqrt(((xy[1:] - xy[:-1]) ** 2).sum(1))).sum()
        assert direct == calc                                  # ... and equals ceil(euclid) on the coordinates
        rs = np.random.RandomState(3)
        cands = random_candidates(rs, vns, T, 200000)
        outs = {}
        for int_path in (True, False):
            e.set_int_path(int_path)
            for seg in ((True, False) if int_path else (False,)):
                e.set_segment_cache(seg)
                outs[(int_path, seg)] = p.eval_batch(p.make_cands(cands))
        e.set_int_path(True)
        e.set_segment_cache(True)
        # the endpoint-vector form (this batch is below its break-even at 2000 clusters: forced), whose per-state build takes
        # its large-instance branches here: int32 row copies, lower-bound sums in a kernel of their own, two-level segment build
        for vec in (2, 3):
            e.set_vector_cache(vec)
            outs[("vec", vec)] = p.eval_batch(p.make_cands(cands))
        e.set_vector_cache(True)
        ln, lb, fl, first = outs[(True, True)]
        for k, (l2, b2, f2, first2) in outs.items():
            assert (l2 == ln).all() and (b2 == lb).all() and (f2 == fl).all() and first2 == first, k
        assert (lb <= ln).all()
        # tryToRemove(i) equals the length after removePoint(i), for a few i
        rem = cands[cands[:, 0] == 2][:3]
        for op, i, _ in rem:
            q = e.path(vns, T)
            want = ln[np.where((cands == (op, i, 0)).all(1))[0][0]]
            q.removePoint(int(i))
            assert q.getPathLength() == want
    finally:
        e.close()


@pytest.mark.gpu
@pytest.mark.parametrize("place,blocks", [(0, 1), (1, 5), (3, 3)])
def test_synthetic_local_search_placements(synth, place, blocks, monkeypatch):
    """The second chain form with its F / B rows in per-block global memory (what instances beyond shared memory use)
    and in shared memory, on 1, 3 and 5 cooperating thread blocks, product tree on: same states as the oracle."""
    from oracle import binding as ob
    from vns_sop_b200 import engine
    e, o, pb, (vns, T) = synth
    monkeypatch.setenv("VNSGPU_CHAIN_PLACE", str(place))
    e.set_chain_blocks(blocks)
    try:
        pg, po = e.path(vns, T), o.path(vns, T)
        pg.fitDOPtoBudget()
        po.fit()
        rg, ro = engine.seeded_rng(5), ob.new_rng(5)
        for k in (1, 2):
            pg.shake(k, rg)
            po.shake(k, ro)
            st = pg._ls(k, 6000, rg, None, 1)
            po.local_search(k, 6000, ro)
            assert st.proposals == 6000
            assert (pg.vns == po.vns).all() and pg.getNumTargets() == po.T
            assert pg.getPathLength() == po.length() and pg.getReward() == po.reward()
        assert engine.rng_next(rg) == ob.rng_next(ro)
    finally:
        e.set_chain_blocks(0)


@pytest.mark.gpu
@pytest.mark.parametrize("staged", [0, 1])
def test_20k_instance_local_search_vs_oracle(staged, monkeypatch):
    """Config 5 (20 000 nodes / 2000 clusters, 3.2 GB table): one insertRandom and one exchangeRandom call of 30 000
    proposals each on the device (cooperating thread blocks, product tree, rows in global memory) against the C port on
    the same state: same VNS vector, targets, length, reward and RNG position; and 4000 candidates of that state
    against the port's tryTo*."""
    import time
    from oracle import binding as ob
    from vns_sop_b200 import engine, instances as ins, synthetic as sy
    monkeypatch.setenv("VNSGPU_VEC_STAGED", str(staged))      # 1: the bulk-copy staged evaluation kernel (opt-in), read at vnsgpu_create
    sf = sy.generate()
    pb = ins.build_problem(sf)
    e = engine.Engine(0).load_problem(pb)
    try:
        o = ob.Oracle(pb)
        vns, T = sy.centroid_tour_state(pb)
        pg, po = e.path(vns, T), o.path(vns, T)
        assert pg.getPathLength() == po.length()
        rs = np.random.RandomState(9)
        cands = random_candidates(rs, vns, T, 4000)
        oln, olb = po.eval_batch(cands)
        for vecmode in (1, 2, 3):      # 1: the library's choice at this batch size (segment products); 2: endpoint vectors forced -
            e.set_vector_cache(vecmode)      # packed vector build + staged evaluation kernel; 3: unpacked vector build
            ln, lb, fl, first = pg.eval_batch(pg.make_cands(cands))
            assert (ln == oln).all() and (lb == olb).all(), vecmode
        e.set_vector_cache(1)
        rg, ro = engine.seeded_rng(3), ob.new_rng(3)
        for k in (1, 2):
            e.sync()
            t0 = time.perf_counter()
            st = pg._ls(k, 30000, rg, None, 1)
            tg = time.perf_counter() - t0
            t0 = time.perf_counter()
            po.local_search(k, 30000, ro)
            tc = time.perf_counter() - t0
            print("20k-node local search kind %d: 30000 proposals, %d accepts, %d rounds: GPU %.3f s, C port (1 core) %.3f s"
                  % (k, st.accepts, st.rounds, tg, tc))
            assert st.proposals == 30000
            assert (pg.vns == po.vns).all() and pg.getNumTargets() == po.T
            assert pg.getPathLength() == po.length() and pg.getReward() == po.reward()
        assert engine.rng_next(rg) == ob.rng_next(ro)
    finally:
        e.close()


def test_dop_generator_is_deterministic():
    from vns_sop_b200 import synthetic as sy
    a, b = sy.generate_dop(120, 8), sy.generate_dop(120, 8)
    assert (a.centers == b.centers).all() and a.budget == b.budget
    assert a.C == 122 and a.N == 122 * 8 and (a.m == 8).all() and a.start == 0 and a.goal == 121
    assert (a.centers[0] == a.centers[-1]).all()               # goal = the depot again (duplicated end set)
    assert a.reward[0] == a.reward[-1] == 0 and (a.reward[1:-1] == 1).all()
    vns, T = sy.centroid_tour_state(a, 0.6)
    assert sorted(vns.tolist()) == list(range(1, 121)) and 5 < T < 120
    big = sy.generate_dop()                                    # the bench workload: 2000 x 32, no table on the host
    assert big.N == 64064 and big.N ** 2 < 2 ** 32 and big.W is None and big.xy is None


@pytest.mark.gpu
def test_synthetic_dop_candidates_vs_oracle():
    """DOP flavour (asymmetric Dubins table built on the device, every cluster 8 heading samples): the table
    and all three evaluation forms against the oracle, which computes the same table on the CPU."""
    from oracle import binding as ob
    from vns_sop_b200 import engine, synthetic as sy
    pb = sy.generate_dop(120, 8)
    e = engine.Engine(0).load_problem(pb, dubins=(8, pb.dubins_radius, 1))
    try:
        o = ob.Oracle(pb, dubins=(8, pb.dubins_radius, 1))
        Wg, Wo = e.table_dense(), o.dense()
        off_diag = ~np.eye(pb.N, dtype=bool)                   # identical states are ill-conditioned
        assert (Wg[off_diag] == Wo[off_diag]).mean() > 0.9999
        pb.W = Wg                                              # evaluate both sides on the device-built table
        o = ob.Oracle(pb)
        vns, T = sy.centroid_tour_state(pb, 0.6)
        pg, po = e.path(vns, T), o.path(vns, T)
        assert pg.getPathLength() == po.length() and pg.getPathLengthRear() == po.length_rear()
        rs = np.random.RandomState(9)
        cands = random_candidates(rs, vns, T, 70000)           # >= C^2 / 4: the endpoint-vector form engages
        oln, olb = po.eval_batch(cands)
        for int_path in (True, False):
            e.set_int_path(int_path)
            for seg, vec in ((True, True), (True, False), (False, False)):
                e.set_segment_cache(seg)
                e.set_vector_cache(vec)
                ln, lb, fl, first = pg.eval_batch(pg.make_cands(cands))
                assert (ln == oln).all() and (lb == olb).all(), (int_path, seg, vec)
    finally:
        e.close()
