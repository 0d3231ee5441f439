"""GPU parity (run on the B200 box): every row of SURVEY.md 8a through the C ABI (ctypes ->
libvnsgpu.so -> CUDA kernels) against (1) golden vectors produced by the reference's own code and
(2) the CPU oracle on the same seeded inputs.  Integer tables: bit-exact.  Real-valued tables:
bit-exact as well (same association, no FMA), asserted with == ; Dubins lengths: 1e-9 relative
(device libm differs from glibc)."""
import numpy as np
import pytest

from conftest import ALL, SMALL, load_arrays, load_problem, random_candidates

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng_cache():
    from vns_sop_b200 import engine
    cache = {}

    def get(name):
        if name not in cache:
            sf, pb = load_problem(name)
            cache[name] = (engine.Engine(0).load_problem(pb), sf, pb)
        return cache[name]
    yield get
    for e, _, _ in cache.values():
        e.close()


def layer_cluster(pb, vns, T, l):
    return pb.start if l == 0 else (pb.goal if l == T + 1 else vns[l - 1])


@pytest.mark.parametrize("name", ALL)
def test_table_and_bounds(name, golden, eng_cache):
    e, sf, pb = eng_cache(name)
    g, a = golden[name], load_arrays(name)
    W = e.table_dense()
    if "dense" in a:
        assert (W == a["dense"]).all()
    else:
        assert float(W.sum()) == g["dense_sum"]
        wsum = (W * (np.arange(pb.N)[:, None] % 7 + 1) * (np.arange(pb.N)[None, :] % 5 + 1)).sum()
        assert float(wsum) == g["dense_wsum"]
    lo, hi = e.bounds()
    assert (lo == a["dmin"]).all() and (hi == a["dmax"]).all()


@pytest.mark.parametrize("rows", ["scan", "products"])
@pytest.mark.parametrize("name", ALL)
def test_update_and_queries_on_greedy_path(name, rows, golden, eng_cache, monkeypatch):
    """update() + the const queries on the reference's greedy tour; rows = how vnsgpu_path_set fills shortest / shortest_back:
    the sequential scan, or every layer in parallel from the exact segment products (the default from 128 targets on)"""
    monkeypatch.setenv("VNSGPU_APPLY_PRODUCTS", "1" if rows == "products" else "0")
    e, sf, pb = eng_cache(name)
    g, a = golden[name], load_arrays(name)
    T = g["greedy"]["T"]
    p = e.path(a["greedy_vns"], T)
    assert p.getNumTargets() == T and (p.getAllTargets() == a["greedy_vns"][:T]).all()
    assert p.getPathLength() == g["greedy"]["length"] and p.getPathLengthRear() == g["greedy"]["length_rear"]
    assert p.getReward() == g["greedy"]["reward"]
    F, B = p.layers()
    for l in range(T + 2):
        k = pb.m[layer_cluster(pb, a["greedy_vns"], T, l)]
        assert (F[l, :k] == a["greedy_F"][l, :k]).all(), l
        assert (B[l, :k] == a["greedy_B"][l, :k]).all(), l
    cl, nd, calc = p.getPath()
    assert (cl == a["greedy_path_cluster"]).all() and (nd == a["greedy_path_node"]).all()
    assert calc == g["greedy"]["length_calc"]
    ln, lb, fl, first = p.eval_batch(a["cands"])
    assert (ln == a["cand_len"]).all()
    assert (lb == a["cand_lb"]).all()
    # flags restated from the reference's predicates (vns_sop_solver.cpp:539,619,784,805)
    L, budget = g["greedy"]["length"], pb.budget
    c = a["cands"]
    dR = np.zeros(len(c))
    dR[c[:, 0] == 0] = pb.reward[c[c[:, 0] == 0, 1]]
    rep = c[:, 0] == 1
    dR[rep] = pb.reward[c[rep, 1]] - pb.reward[a["greedy_vns"][c[rep, 2]]]
    feas = ln <= budget
    acc = feas & ((dR > 0) | ((L - ln) > 1e-5))
    skip = (lb > budget) & (c[:, 0] != 2)
    assert ((fl & 1) != 0).tolist() == feas.tolist()
    assert ((fl & 2) != 0).tolist() == acc.tolist()
    assert ((fl & 4) != 0).tolist() == skip.tolist()
    ok = np.nonzero(acc & ~skip)[0]
    assert first == (int(ok[0]) if len(ok) else -1)
    # single-call mirrors of the reference methods
    k = int(np.nonzero(c[:, 0] == 3)[0][0])
    assert p.tryToMove(c[k, 1], c[k, 2]) == a["cand_len"][k]
    assert p.tryToMoveLowerBound(c[k, 1], c[k, 2]) == a["cand_lb"][k]


@pytest.mark.parametrize("name", SMALL)
def test_add_remove_point_incremental(name, golden, eng_cache):
    from oracle import binding as ob
    e, sf, pb = eng_cache(name)
    a = load_arrays(name)
    o = ob.Oracle(pb)
    po, pg = o.path(), e.path()
    rs = np.random.RandomState(3)
    inside = []
    outside = list(a["initial_vns"])
    for step in range(40):
        if inside and (rs.rand() < 0.35 or not outside):
            idx = int(rs.randint(0, len(inside)))
            outside.append(inside.pop(idx))
            po.remove_point(idx)
            pg.removePoint(idx)
        else:
            c = outside.pop(int(rs.randint(0, len(outside))))
            idx = int(rs.randint(0, len(inside) + 1))
            inside.insert(idx, c)
            po.add_point(c, idx)
            pg.addPoint(c, idx)
        assert pg.getNumTargets() == po.T and (pg.getAllTargets() == po.targets).all()
        assert pg.getPathLength() == po.length() and pg.getPathLengthRear() == po.length_rear()
        Fo, Bo, _, _ = po.layers()
        F, B = pg.layers()
        for l in range(po.T + 2):
            k = pb.m[layer_cluster(pb, inside, po.T, l)]
            assert (F[l, :k] == Fo[l, :k]).all() and (B[l, :k] == Bo[l, :k]).all()


@pytest.mark.parametrize("name", ALL)
def test_greedy_insertion(name, golden, eng_cache):
    e, sf, pb = eng_cache(name)
    g, a = golden[name], load_arrays(name)
    p = e.path()
    p.greedy_insertion()
    assert p.getNumTargets() == g["greedy"]["T"] and (p.vns == a["greedy_vns"]).all()
    assert p.getPathLength() == g["greedy"]["length"] and p.getReward() == g["greedy"]["reward"]


CHAIN_MODES = ["smem-tab", "smem-tab-u16", "smem-tab-g32", "smem-tab-g16", "smem-int", "smem-int-tree", "smem-int-grid3", "smem-int-grid8-tree", "smem-int-v1", "smem-double", "global"]


def _set_chain_mode(e, mode):
    """smem-int: second chain form (lower bound first), one thread block; -gridN: N cooperating thread blocks;
    -tree: with the product tree forced on (default: off below 96 clusters); -v1: first form; smem-double: first form
    in IEEE double; global: global-memory chain; smem-tab: third form (the whole distance table in shared memory, as
    int32 when that fits, else as uint16), -u16: its uint16 copy forced (instances without one take another form), -g32 /
    -g16: the table read from global memory (int32 / uint16 copy), state in shared memory"""
    import os
    e.set_smem_chain(mode != "global")
    e.set_int_path(mode != "smem-double")
    e.set_chain_form(3 if mode.startswith("smem-tab") else (1 if mode == "smem-int-v1" else 0))
    os.environ.pop("VNSGPU_CHAIN3_U16", None)
    if mode in ("smem-tab-u16", "smem-tab-g32", "smem-tab-g16"):      # variant: bit 0 uint16 copy, bit 1 table read from global memory
        os.environ["VNSGPU_CHAIN3_U16"] = {"smem-tab-u16": "1", "smem-tab-g32": "2", "smem-tab-g16": "3"}[mode]
    e.set_chain_blocks(int(mode.split("grid")[1].split("-")[0]) if "grid" in mode else 1)
    e.set_chain_tree(1 if "tree" in mode else (0 if mode.startswith("smem-int") else -1))


def _reset_chain_mode(e):
    import os
    os.environ.pop("VNSGPU_CHAIN3_U16", None)
    e.set_smem_chain(True)
    e.set_int_path(True)
    e.set_chain_form(-1)
    e.set_chain_blocks(0)
    e.set_chain_tree(-1)


@pytest.mark.parametrize("mode", CHAIN_MODES)
@pytest.mark.parametrize("name", ALL)
def test_shake_and_local_search_sequence(name, mode, golden, eng_cache):
    """pathInsert/pathExchange + insertRandom/exchangeRandom with seed 7: the device chain must
    walk the reference's exact state sequence and leave the RNG at the same position.  Every
    implementation: second chain form on 1, 3 and 8 cooperating thread blocks, first form in exact int32 and in
    double arithmetic, global-memory chain."""
    e, sf, pb = eng_cache(name)
    g, a = golden[name], load_arrays(name)
    _set_chain_mode(e, mode)
    try:
        _run_sequence(e, pb, g, a)
    finally:
        _reset_chain_mode(e)


@pytest.mark.parametrize("mode", ["smem-tab", "smem-tab-g16", "smem-int", "smem-int-v1"])
@pytest.mark.parametrize("name", ["berlin52", "opn_s08", "dop_s12", "krob200"])
def test_local_search_edge_cases_vs_oracle(name, mode, golden, eng_cache):
    """insertRandom / exchangeRandom calls the golden sequence does not reach, against the C port on the same state and
    generator: calls of 1 .. 100 proposals (the window's cut at the end of a call), the lower-bound skip off, a budget nothing
    fits, a budget everything fits, an empty tour.  States, counters and generator positions must agree after every call."""
    from oracle import binding as ob
    from vns_sop_b200 import engine
    e, sf, pb = eng_cache(name)
    g, a = golden[name], load_arrays(name)
    o = ob.Oracle(pb)
    _set_chain_mode(e, mode)
    try:
        for start in ("greedy", "empty"):
            vns = np.array(a["greedy_vns"])
            T = g["greedy"]["T"] if start == "greedy" else 0
            pg, po = e.path(vns, T), o.path(vns, T)
            rg, ro = engine.seeded_rng(11), ob.new_rng(11)
            cases = [(1, 1, None, 1), (2, 1, None, 1), (1, 5, None, 1), (2, 33, None, 0), (1, 100, None, 0), (2, 64, None, 1),
                     (1, 40, 1.0, 1), (2, 40, 1.0, 0), (1, 70, 1e12, 1), (2, 70, 1e12, 1), (1, 31, pb.budget * 0.5, 1), (2, 32, pb.budget * 0.5, 0)]
            for i, (kind, iters, budget, skip) in enumerate(cases):
                sg = pg._ls(kind, iters, rg, budget, skip)
                so = po.local_search(kind, iters, ro, budget, skip)
                assert pg.getNumTargets() == po.T and (pg.vns == po.vns).all(), (start, i)
                assert pg.getPathLength() == po.length() and pg.getReward() == po.reward(), (start, i)
                assert (sg.proposals, sg.accepts) == (so.proposals, so.accepts), (start, i)
                assert sg.lb_skips == so.lb_skips and sg.evals == so.evals, (start, i)
            assert engine.rng_next(rg) == ob.rng_next(ro), start
    finally:
        _reset_chain_mode(e)


def _run_sequence(e, pb, g, a):
    from vns_sop_b200 import engine
    p = e.path(a["greedy_vns"], g["greedy"]["T"])
    rng = engine.seeded_rng(7)
    for i, st in enumerate(g["sequence"]):
        if st["op"] == "shake":
            p.shake(st["k"], rng)
        else:
            iters = len(p.vns) ** 2 if g["sequence_ls_iters"] == "nvns^2" else g["sequence_ls_iters"]
            stats = p._ls(st["k"], iters, rng, None, 1)
            assert stats.proposals == iters
        assert p.getNumTargets() == st["T"], (i, st)
        assert p.getPathLength() == st["length"] and p.getReward() == st["reward"], (i, st)
        assert (p.vns == a["sequence_vns"][i]).all(), (i, st)
        assert (p.getAllTargets() == a["sequence_vns"][i][:st["T"]]).all()
    assert engine.rng_next(rng) == g["rand_after_sequence"]


@pytest.mark.parametrize("mode", CHAIN_MODES)
@pytest.mark.parametrize("name", ["berlin52", "opn_s08", "dop_s12", "krob200"])
def test_full_vns_run_seed1_follows_reference_trajectory(name, mode, golden, eng_cache):
    """VnsSopSolver::iterate with the stock stopping rule and seed 1 as one device chain: iteration
    counts, every improvement (reward, length) and the final tour equal the reference's results.log."""
    e, sf, pb = eng_cache(name)
    s = golden[name]["solve_seed1"]
    if mode not in ("smem-tab", "smem-int", "smem-int-tree", "smem-int-grid3") and name == "krob200":
        pytest.skip("long run: the third and the second chain form (1 and 3 thread blocks) cover krob200")
    _set_chain_mode(e, mode)
    try:
        p = e.path()
        res, impr = p.vns_run(1)
    finally:
        _reset_chain_mode(e)
    assert res.status == 0
    assert res.num_iters == s["NUM_ITERS"] and res.num_iters_last_impr == s["NUM_ITERS_LAST_IMPR"]
    assert res.reward == s["REWARDS"] and res.length == s["LENGTH"]
    assert [[r, l] for r, l, _ in impr] == s["IMPROVEMENTS"]
    cl, nd, calc = p.getPath()
    assert list(pb.node_id[pb.off[cl] + nd]) == s["RESULT_TARGET_IDS"]
    assert list(pb.orig_cluster_id[cl]) == s["RESULT_CLUSTER_IDS"]
    assert calc <= pb.budget


def test_multi_start_batch_equals_single_seeded_runs(golden, eng_cache):
    """vnsgpu_vns_run_batch (the reference's --batch=N as one launch, one thread block per seed): every run
    is the same trajectory as the single-seed call, seed 1 is the reference's results.log, and the best
    tour lands in the path."""
    for name in ("berlin52", "opn_s08"):
        e, sf, pb = eng_cache(name)
        seeds = [1, 2, 3, 4, 5, 6, 7]
        p = e.path()
        res, best = p.vns_run_batch(seeds)
        s = golden[name]["solve_seed1"]
        assert res[0].reward == s["REWARDS"] and res[0].length == s["LENGTH"] and res[0].num_iters == s["NUM_ITERS"]
        for k, sd in enumerate(seeds):
            q = e.path()
            r1, _ = q.vns_run(sd)
            assert (res[k].reward, res[k].length, res[k].num_iters, res[k].num_iters_last_impr, res[k].proposals, res[k].accepts) == \
                   (r1.reward, r1.length, r1.num_iters, r1.num_iters_last_impr, r1.proposals, r1.accepts), (name, sd)
        key = [(-r.reward, r.length) for r in res]
        assert best == key.index(min(key))
        assert p.getReward() == res[best].reward and p.getPathLength() == res[best].length
        assert p.getPathLengthCalculate() <= pb.budget


def test_multi_start_race_stops_at_target_reward(golden, eng_cache):
    """vnsgpu_set_target_reward: the launch ends once one seed's tour reaches the target; every run is then a PREFIX
    of its seeded trajectory (its tour is one of the improvements the full run logs), the winner holds >= target,
    and with the target switched off again the batch is the full runs."""
    for name in ("berlin52", "opn_s08"):
        e, sf, pb = eng_cache(name)
        seeds = [1, 2, 3, 4, 5, 6, 7, 8]
        full = {}
        for sd in seeds:
            r1, impr = e.path().vns_run(sd)
            full[sd] = (r1, [(r, l) for r, l, _ in impr])
        target = sorted(r.reward for r, _ in full.values())[len(seeds) // 2]       # the median final reward
        p = e.path()
        res, best = p.vns_run_batch(seeds, target_reward=target)
        assert res[best].reward >= target and p.getReward() == res[best].reward
        assert any(r.num_iters < full[sd][0].num_iters for r, sd in zip(res, seeds)), "nobody stopped early"
        for r, sd in zip(res, seeds):
            assert r.status == 0 and r.num_iters <= full[sd][0].num_iters
            assert (r.reward, r.length) in full[sd][1], (name, sd)
        res2, _ = e.path().vns_run_batch(seeds)                              # target off again (0)
        for r, sd in zip(res2, seeds):
            assert (r.reward, r.length, r.num_iters) == (full[sd][0].reward, full[sd][0].length, full[sd][0].num_iters)


@pytest.mark.parametrize("gx", [1, 3])
@pytest.mark.parametrize("name", ["krob200", "dop_s12", "vm1084"])
def test_endpoint_vectors_do_not_depend_on_the_grid(name, gx, golden, eng_cache, monkeypatch):
    """k_vec_pairs strides over the layers with the warps of grid.x, which the host sizes from the header as it last read it:
    a grid far below the layers (symmetric and asymmetric tables) gives the oracle's lengths"""
    from oracle import binding as ob
    e, sf, pb = eng_cache(name)
    g, a = golden[name], load_arrays(name)
    monkeypatch.setenv("VNSGPU_VEC_PAIRS_GX", str(gx))
    monkeypatch.setenv("VNSGPU_SEG_CROSS_GRID", str(gx))      # the same for the cross products of the two-level segment build
    po, pg = ob.Oracle(pb).path(a["greedy_vns"], g["greedy"]["T"]), e.path(a["greedy_vns"], g["greedy"]["T"])
    cands = random_candidates(np.random.RandomState(11), po.vns, po.T, 300007 if name == "vm1084" else 60013)
    oln, olb = po.eval_batch(cands)
    try:
        for seg in (1, 3):                                    # 3: two-level build whatever the size
            e.set_segment_cache(seg)
            ln, lb, fl, first = pg.eval_batch(cands)
            assert (ln == oln).all() and (lb == olb).all(), seg
    finally:
        e.set_segment_cache(True)


@pytest.mark.parametrize("int_path", [True, False])
@pytest.mark.parametrize("name", ALL)
def test_eval_batch_both_arithmetic_paths_vs_oracle(name, int_path, golden, eng_cache):
    """The sub-warp group kernel in exact int32 (DPX) and in IEEE double arithmetic: lengths, lower
    bounds and the first accepted index against the CPU oracle on several search states, with
    out-of-range candidates mixed in and a batch size that leaves groups and warps ragged."""
    from oracle import binding as ob
    from vns_sop_b200 import engine
    e, sf, pb = eng_cache(name)
    g, a = golden[name], load_arrays(name)
    e.set_int_path(int_path)
    try:
        assert e.int_path_active == int_path        # every shipped instance has an all-integer table
        o = ob.Oracle(pb)
        po, pg = o.path(a["greedy_vns"], g["greedy"]["T"]), e.path(a["greedy_vns"], g["greedy"]["T"])
        go, gg = ob.new_rng(21), engine.seeded_rng(21)
        rs = np.random.RandomState(5)
        for it in range(4):
            T = po.T
            if T >= 3:
                cands = random_candidates(rs, po.vns, T, 2999)
                cands[::97] = [3, 0, 1]                 # |what - where| < 2: the reference exit(1)s
                ln, lb, fl, first = pg.eval_batch(cands)
                oln, olb = po.eval_batch(cands)
                bad = (fl & engine.F_BAD) != 0
                assert bad[::97].all() and bad.sum() == len(cands[::97])
                assert (ln[~bad] == oln[~bad]).all() and (lb[~bad] == olb[~bad]).all()
                ok = np.nonzero(((fl & 2) != 0) & ((fl & 4) == 0))[0]
                assert first == (int(ok[0]) if len(ok) else -1)
            k = 1 + it % 2
            po.shake(k, go)
            pg.shake(k, gg)
            po.local_search(k, 400, go)
            pg._ls(k, 400, gg, None, 1)
            assert (pg.vns == po.vns).all() and pg.getNumTargets() == po.T
        # a batch large enough for the chain-length counting sort (and, through the host-buffer
        # call, for several pipeline chunks); same answers with the sort switched off
        T = po.T
        if T >= 3:
            big = random_candidates(rs, po.vns, T, {"krob200": 150001, "vm1084": 200003}.get(name, 20011))
            big[::1013] = [4, 0, 0]
            ln, lb, fl, first = pg.eval_batch(big)
            oln, olb = po.eval_batch(big)
            bad = (fl & engine.F_BAD) != 0
            assert bad.sum() == len(big[::1013])
            assert (ln[~bad] == oln[~bad]).all() and (lb[~bad] == olb[~bad]).all()
            ok = np.nonzero(((fl & 2) != 0) & ((fl & 4) == 0))[0]
            assert first == (int(ok[0]) if len(ok) else -1)
            import torch, ctypes as C                                   # sharded device entry point: global key
            d_c = torch.from_numpy(pg.make_cands(big).view(np.uint8).copy()).cuda()
            d_key = torch.zeros(1, dtype=torch.int64, device="cuda")
            for lo_, hi_ in ((0, len(big)), (len(big) // 2, len(big))):
                e._ck(e.lib.vnsgpu_eval_batch_dev_shard(e.h, pg.h, hi_ - lo_, C.c_void_p(d_c.data_ptr() + 16 * lo_), pb.budget, 1,
                                                        None, None, None, C.c_void_p(d_key.data_ptr()), 1000 + lo_, None))
                e.sync()
                okk = ok[ok >= lo_]
                assert int(d_key.item()) == (1000 + int(okk[0]) if len(okk) else engine.NO_KEY)
            lnp, lbp, flp, firstp = pg.eval_batch_packed(big)          # 8-byte descriptors, same answers
            assert (lnp[~bad] == ln[~bad]).all() and (lbp[~bad] == lb[~bad]).all() and (flp == fl).all() and firstp == first
            if int_path:     # compact wire format: 4-byte descriptors in, int32 lengths out (exact on integer tables)
                ln32, fl32, first32 = pg.eval_batch_packed32(big)
                assert ln32.dtype == np.int32 and (ln32[~bad] == ln[~bad]).all() and (ln32[bad] == np.iinfo(np.int32).min).all()
                assert (fl32 == fl).all() and first32 == first
                lnw, flw, firstw = pg.eval_batch_packed32r(big)      # one result word per candidate: length | flags << 29
                assert (lnw == ln32).all() and (flw == fl).all() and firstw == first
            else:            # real-valued evaluation path: the int32 call refuses instead of rounding
                with pytest.raises(engine.VnsGpuError):
                    pg.eval_batch_packed32(big[:64])
            # (the integer path has just used the endpoint vectors + exact segment products: batch >= 16384
            # and >= C^2 / 4); the same batch with the vectors off, the products off, the sort off
            for sort, seg, vec in ((False, True, True), (True, True, False), (False, True, False), (True, False, False), (False, False, False)):
                e.set_batch_sort(sort)
                e.set_segment_cache(seg)
                e.set_vector_cache(vec)
                ln2, lb2, fl2, first2 = pg.eval_batch(big)
                assert (ln2[~bad] == ln[~bad]).all() and (lb2[~bad] == lb[~bad]).all() and (fl2 == fl).all()
                assert first2 == first
                if int_path and vec:         # the endpoint-vector builds from the int32 table instead of the uint16 copy
                    e.set_u16_table(False)
                    ln3, lb3, fl3, first3 = pg.eval_batch(big)
                    e.set_u16_table(True)
                    assert (ln3[~bad] == ln[~bad]).all() and (lb3[~bad] == lb[~bad]).all() and (fl3 == fl).all() and first3 == first
                if int_path and not vec:     # result words through the conversion kernel (sorted / block-by-block forms)
                    lnw, flw, firstw = pg.eval_batch_packed32r(big)
                    assert (lnw[~bad] == ln[~bad]).all() and (flw == fl).all() and firstw == first
    finally:
        e.set_int_path(True)
        e.set_batch_sort(True)
        e.set_segment_cache(True)
        e.set_vector_cache(True)


@pytest.mark.parametrize("name", ["krob200", "opn_s08"])
def test_host_call_graph_replay(name, golden, eng_cache):
    """The host-buffer call with PINNED buffers becomes one CUDA graph launch from the third identical call on.
    The graph must stay valid while the path's state changes on the device (the kernels read targets, T, F / B
    there) and must be dropped when a buffer or the batch size changes: every call is compared with the
    eager call on pageable buffers (bit-exact) through several local-search steps."""
    import ctypes as C
    import torch
    from vns_sop_b200 import engine
    e, sf, pb = eng_cache(name)
    g, a = golden[name], load_arrays(name)
    pg = e.path(a["greedy_vns"], g["greedy"]["T"])
    rs = np.random.RandomState(11)
    T = pg.getNumTargets()
    base = random_candidates(rs, pg.vns, T, 20000)
    cands = np.tile(base, (15, 1))                      # 300 000 >= the graph threshold
    rs.shuffle(cands)
    n = len(cands)
    pk = torch.from_numpy(pg.pack_cands32(cands).view(np.int32).copy()).pin_memory()
    h_len = torch.empty(n, dtype=torch.int32).pin_memory()
    h_len2 = torch.empty(n, dtype=torch.int32).pin_memory()
    h_fl = torch.empty(n, dtype=torch.uint8).pin_memory()
    first = C.c_int64(-1)
    rng = engine.seeded_rng(5)

    def pinned_call(nn, out):
        e._ck(e.lib.vnsgpu_eval_batch_packed32(e.h, pg.h, nn, C.c_void_p(pk.data_ptr()), pb.budget, 1,
                                               C.c_void_p(out.data_ptr()), C.c_void_p(h_fl.data_ptr()), C.byref(first)))
        return out.numpy()[:nn].copy(), h_fl.numpy()[:nn].copy(), first.value

    l0 = e.launches
    for it in range(6):
        ln, fl, f1 = pinned_call(n, h_len)               # calls 0, 1 eager; 2.. replay the graph
        rln, rlb, rfl, rf = pg.eval_batch(cands)         # pageable buffers: always eager
        bad = (rfl & engine.F_BAD) != 0
        assert (ln[~bad] == rln[~bad]).all() and (fl == rfl).all() and f1 == rf, (name, it)
        assert (ln[bad] == np.iinfo(np.int32).min).all()
        pg._ls(1 + it % 2, 40, rng, None, 1)             # the search moves on: T, targets, F / B change on the device
    assert e.launches > l0                               # replays still count their kernels
    # another output buffer, another batch size: the recorded graph must not be reused
    ln, fl, f1 = pinned_call(n, h_len2)
    rln, rlb, rfl, rf = pg.eval_batch(cands)
    bad = (rfl & engine.F_BAD) != 0
    assert (ln[~bad] == rln[~bad]).all() and (fl == rfl).all() and f1 == rf
    m = n - 4099
    for _ in range(4):
        ln, fl, f1 = pinned_call(m, h_len2)
    rln, rlb, rfl, rf = pg.eval_batch(cands[:m])
    bad = (rfl & engine.F_BAD) != 0
    assert (ln[~bad] == rln[~bad]).all() and (fl == rfl).all() and f1 == rf


@pytest.mark.parametrize("S", [4, 8, 12])
def test_dubins_table_known_answers(S):
    """a4: the Dubins kernel against the 98 784 entries shipped with the reference's DOP datasets
    and against the CPU oracle's lengths."""
    from oracle import binding as ob
    from vns_sop_b200 import engine
    sf, pb = load_problem("dop_s%02d" % S)
    e = engine.Engine(0).load_problem(pb, dubins=(S, sf.dubins_radius, 0))
    L = e.table_dense()
    W = pb.W
    # identical start and end state (the diagonal): the true length is 0, but LSR/RSL evaluate
    # mod2pi(tau - alpha) at tau - alpha = +-1 ulp, so any libm returns 0 or one full circle
    # (the shipped s_12 file itself holds 315 = ceil(2 pi rho) in 21 of its 252 diagonal
    # entries).  No solver path reads a same-cluster block.
    diag = np.eye(pb.N, dtype=bool)
    full = 2 * np.pi * sf.dubins_radius
    dl = L[diag]
    assert (np.minimum(np.abs(dl), np.abs(dl - full)) < 1e-9).all()
    # knife-edge entries (a CSC word with a zero-length straight segment, kept by the kernel and the
    # oracle, kept or dropped by the files' generator): see tests/test_oracle_golden.py
    Lo = ob.Oracle(pb, dubins=(S, sf.dubins_radius, 0)).dense()
    Ls = ob.Oracle(pb, dubins=(S, sf.dubins_radius, 0, -1e-9)).dense()
    knife = (Lo != Ls) & ~diag

    def kat(X):
        return (np.ceil(X) == W) | (np.ceil(X - 1e-9) == W) | (np.ceil(X + 1e-9) == W)
    ok = kat(L) | (knife & kat(Ls)) | diag
    bad = np.argwhere(~ok)
    assert len(bad) == 0, [(int(i), int(j), float(L[i, j]), float(W[i, j])) for i, j in bad[:8]]
    close = np.isclose(L, Lo, rtol=1e-9, atol=1e-9) | diag
    bad = np.argwhere(~close)
    assert len(bad) == 0, [(int(i), int(j), float(L[i, j]), float(Lo[i, j])) for i, j in bad[:8]]
    e2 = engine.Engine(0).load_problem(pb, dubins=(S, sf.dubins_radius, 1))
    Lc = e2.table_dense()
    assert (Lc == np.ceil(L)).all()
    near = (np.abs(L - np.round(L)) < 1e-9) | diag | knife
    assert (Lc[~near] == W[~near]).all()
    e.close()
    e2.close()


@pytest.mark.parametrize("rule", [0, 1])
def test_real_valued_and_rounded_tables_vs_oracle(rule):
    """a3 (raw Euclid, real-valued distances: exercises the floating-point association of
    SURVEY.md App. A.4) and EUC_2D rounding; whole driver sequences against the CPU oracle."""
    import dataclasses
    from oracle import binding as ob
    from vns_sop_b200 import engine
    sf, pb = load_problem("krob200")
    rs = np.random.RandomState(11)
    pb = dataclasses.replace(pb, rule=rule, xy=pb.xy + rs.rand(*pb.xy.shape))
    e = engine.Engine(0).load_problem(pb)
    o = ob.Oracle(pb)
    assert (e.table_dense() == o.dense()).all()
    lo, hi = e.bounds()
    olo, ohi = o.bounds()
    assert (lo == olo).all() and (hi == ohi).all()
    po, pg = o.path(), e.path()
    po.greedy()
    pg.greedy_insertion()
    assert (pg.vns == po.vns).all() and pg.getPathLength() == po.length()
    go, gg = ob.new_rng(5), engine.seeded_rng(5)
    for it in range(6):
        k = 1 + it % 2
        po.shake(k, go)
        pg.shake(k, gg)
        assert (pg.vns == po.vns).all() and pg.getNumTargets() == po.T and pg.getPathLength() == po.length()
        if po.T >= 3:
            cands = random_candidates(rs, po.vns, po.T, 300)
            ln, lb, fl, _ = pg.eval_batch(cands)
            oln, olb = po.eval_batch(cands)
            assert (ln == oln).all() and (lb == olb).all()
        po.local_search(k, 1600, go)
        pg._ls(k, 1600, gg, None, 1)
        assert (pg.vns == po.vns).all() and pg.getNumTargets() == po.T
        assert pg.getPathLength() == po.length() and pg.getReward() == po.reward()
    assert engine.rng_next(gg) == ob.rng_next(go)
    e.close()


def test_edge_cases(eng_cache):
    from vns_sop_b200 import engine
    e, sf, pb = eng_cache("berlin52")
    p = e.path()                                   # empty path: start -> goal
    assert p.getNumTargets() == 0 and p.getPathLength() == 0.0   # start and goal are the same set
    ln, lb, fl, first = p.eval_batch(np.zeros((0, 3), np.int32))
    assert len(ln) == 0 and first == -1
    bad = np.array([[3, 0, 1], [4, 0, 0], [0, 999, 0], [2, 5, 0], [1, 3, 0], [7, 0, 0]], np.int32)
    ln, lb, fl, first = p.eval_batch(bad)
    assert (fl == engine.F_BAD).all() and np.isnan(ln).all() and first == -1
    with pytest.raises(engine.VnsGpuError):
        p.removePoint(0)                           # reference: exit(1) (vns_sop_path.cpp:158-161)
    with pytest.raises(engine.VnsGpuError):
        p.addPoint(3, 2)                           # vns_sop_path.cpp:69-72
    with pytest.raises(engine.VnsGpuError):
        p.tryToMove(0, 1)
    # a descriptor whose gain cluster is out of range is a bad descriptor (it indexes the reward table)
    q = e.path()
    q.greedy_insertion()
    c = q.make_cands(np.array([[0, int(q.vns[-1]), 0]], np.int32))
    c["gain_cluster"] = pb.C + 5
    ln, lb, fl, first = q.eval_batch(c)
    assert fl[0] == engine.F_BAD and np.isnan(ln[0]) and first == -1
    # tours are capped at one target per cluster (the batched evaluation is sized for that): duplicates are refused
    # with a status, never a crash
    r = e.path()
    with pytest.raises(engine.VnsGpuError) as ei:
        for k in range(pb.C + 2):
            r.addPoint(2 + k % 3, 0)
    assert "E_RANGE" in str(ei.value) and r.getNumTargets() == pb.C
    # local search on states where the reference's guards skip the loop (:498, :711)
    rng = engine.seeded_rng(1)
    before = engine.rng_next(engine.Rng.from_buffer_copy(rng))
    st = p.exchangeRandom(100, rng)                # T == 0 -> no proposals, RNG untouched
    assert st.proposals == 0 and engine.rng_next(rng) == before


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["krob200", "vm1084"])
def test_device_side_proposals_equal_the_host_draws(name, eng_cache):
    """vnsgpu_gen_proposals_dev (every thread block jumps to its slice of the glibc stream) against vnsgpu_gen_proposals
    (the sequential replay of vns_sop_solver.cpp:507-519, :723-730): same proposals in the same order in both descriptor
    formats, same generator afterwards, and the per-proposal draw counts rewind to the state the sequential loop has
    right after that proposal.  vnsgpu_propose_eval = generate + evaluate + first accepted + rewind."""
    import ctypes as C
    import torch
    from vns_sop_b200 import engine
    e, sf, pb = eng_cache(name)
    p = e.path()
    p.greedy_insertion()
    dev = torch.device("cuda", 0)

    def same_stream(a, b):      # the same generator, whatever the rotation of the 31-word ring
        a, b = engine.Rng.from_buffer_copy(bytes(a)), engine.Rng.from_buffer_copy(bytes(b))
        return [engine.rng_next(a) for _ in range(64)] == [engine.rng_next(b) for _ in range(64)]

    for kind in (1, 2):
        for n in (1000, 300000):
            r_host, r_dev, r_dev2 = engine.seeded_rng(21 + kind), engine.seeded_rng(21 + kind), engine.seeded_rng(21 + kind)
            host = p.gen_proposals(kind, n, r_host)
            d_c = torch.empty(n * 16, dtype=torch.uint8, device=dev)
            d_dr = torch.empty(n, dtype=torch.int32, device=dev)
            p.gen_proposals_dev(kind, n, r_dev, d_c.data_ptr(), 0, d_dr.data_ptr())
            got = d_c.cpu().numpy().view(engine.CAND_DTYPE)
            for f in ("op", "a", "b", "gain_cluster"):
                assert (got[f] == host[f]).all(), (kind, n, f)
            assert same_stream(r_dev, r_host)
            d_w = torch.empty(n, dtype=torch.int32, device=dev)
            p.gen_proposals_dev(kind, n, r_dev2, d_w.data_ptr(), 2, None)
            assert (d_w.cpu().numpy().view(np.uint32) == p.pack_cands32(host)).all()
            assert same_stream(r_dev2, r_host)
            draws = d_dr.cpu().numpy().view(np.uint32)
            assert (np.diff(draws.astype(np.int64)) >= 2).all() and draws[0] >= 2
            for k in (0, 7, n // 2, n - 1):            # generator right after proposal k
                a, b = engine.seeded_rng(21 + kind), engine.seeded_rng(21 + kind)
                p.gen_proposals(kind, k + 1, a)
                engine.rng_advance(b, int(draws[k]))
                assert engine.rng_next(a) == engine.rng_next(b) and engine.rng_next(a) == engine.rng_next(b), (kind, n, k)
        # one round: first accepted proposal of a window + generator rewound to it
        n = 5000
        r1, r2 = engine.seeded_rng(77), engine.seeded_rng(77)
        host = p.gen_proposals(kind, n, engine.seeded_rng(77))
        _, _, fl, first = p.eval_batch(host)
        got_first, cand = p.propose_eval(kind, n, r1)
        assert got_first == first
        if first >= 0:
            assert cand == tuple(int(host[f][first]) for f in ("op", "a", "b", "gain_cluster"))
            p.gen_proposals(kind, first + 1, r2)
        else:
            p.gen_proposals(kind, n, r2)
        assert engine.rng_next(r1) == engine.rng_next(r2)


@pytest.mark.gpu
def test_real_valued_explicit_matrix_double_path():
    """The fixture tests/test_oracle_vs_ref.py pins against the compiled reference (real-valued EXPLICIT matrix), on the
    GPU's IEEE double path: table, bounds, greedy, 20 000 candidates, shake + local search, generator position."""
    from conftest import real_valued_explicit_krob200
    from oracle import binding as ob
    from vns_sop_b200 import engine
    sf, pb = real_valued_explicit_krob200()
    e = engine.Engine(0).load_problem(pb)
    try:
        assert not e.int_path_active
        o = ob.Oracle(pb)
        assert (e.table_dense() == o.dense()).all()
        lo, hi = e.bounds()
        olo, ohi = o.bounds()
        assert (lo == olo).all() and (hi == ohi).all()
        po, pg = o.path(), e.path()
        po.greedy()
        pg.greedy_insertion()
        assert (pg.vns == po.vns).all() and pg.getPathLength() == po.length()
        rs = np.random.RandomState(3)
        cands = random_candidates(rs, po.vns, po.T, 20000)
        ln, lb, fl, _ = pg.eval_batch(cands)
        oln, olb = po.eval_batch(cands)
        assert (ln == oln).all() and (lb == olb).all()
        go, gg = ob.new_rng(9), engine.seeded_rng(9)
        for it in range(10):
            k = 1 + it % 2
            po.shake(k, go)
            pg.shake(k, gg)
            po.local_search(k, 1600, go)
            pg._ls(k, 1600, gg, None, 1)
            assert (pg.vns == po.vns).all() and pg.getNumTargets() == po.T
            assert pg.getPathLength() == po.length() and pg.getReward() == po.reward()
        assert engine.rng_next(gg) == ob.rng_next(go)
    finally:
        e.close()


@pytest.mark.gpu
def test_dop_s16_generated_instance_vs_oracle():
    """BASELINE config 3: 16 heading samples on the Tsiligirides-2 centres, table from the Dubins kernel; the C port
    loads that table as the EXPLICIT matrix the reference's DOP files carry.  Greedy, candidates in every evaluation
    form, a seeded shake + local-search sequence."""
    import bench
    from oracle import binding as ob
    from vns_sop_b200 import engine
    sf, pb = bench.load_workload("dop_s16")
    e = engine.Engine(0).load_problem(pb, dubins=(16, sf.dubins_radius, 1))
    try:
        bench.fill_dubins_weights(sf, pb, e.table_dense())
        o = ob.Oracle(pb)
        assert (e.table_dense() == o.dense()).all() and pb.m.max() == 16
        po, pg = o.path(), e.path()
        po.greedy()
        pg.greedy_insertion()
        assert (pg.vns == po.vns).all() and pg.getPathLength() == po.length() and pg.getReward() == po.reward()
        rs = np.random.RandomState(4)
        cands = random_candidates(rs, po.vns, po.T, 30000)
        oln, olb = po.eval_batch(cands)
        for int_path in (True, False):
            e.set_int_path(int_path)
            for seg, vec in ((1, 2), (1, 0), (0, 0)):
                e.set_segment_cache(seg)
                e.set_vector_cache(vec)
                ln, lb, fl, _ = pg.eval_batch(cands)
                assert (ln == oln).all() and (lb == olb).all(), (int_path, seg, vec)
        e.set_int_path(True); e.set_segment_cache(1); e.set_vector_cache(1)
        go, gg = ob.new_rng(6), engine.seeded_rng(6)
        for it in range(6):
            k = 1 + it % 2
            po.shake(k, go)
            pg.shake(k, gg)
            po.local_search(k, len(po.vns) ** 2, go)
            pg._ls(k, len(po.vns) ** 2, gg, None, 1)
            assert (pg.vns == po.vns).all() and pg.getNumTargets() == po.T and pg.getPathLength() == po.length()
        assert engine.rng_next(gg) == ob.rng_next(go)
    finally:
        e.close()
