"""Full-neighbourhood modes through the C ABI: the batched greedy_insertion must build the
reference's greedy tour (golden vectors from the reference's own code); vnsgpu_best_move must pick, bit for bit, the
candidate a brute-force pass over the oracle's tryTo* values picks under the documented order; the descent must end in
a state without improving candidate."""
import numpy as np
import pytest

from conftest import ALL, SMALL, load_arrays, load_problem

pytestmark = pytest.mark.gpu

OP_ADD, OP_REPLACE, OP_REMOVE, OP_MOVE, OP_EXCHANGE = range(5)
EPS = 0.00001   # MIN_CHANGE_EPS, vns_sop_solver.cpp:15


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


@pytest.fixture(params=["rescan", "products"])
def apply_form(request, monkeypatch):
    """how an accepted move reaches the DP rows: the one-block sequential re-scan, or every layer in parallel from the segment
    products of the new state (the default from 256 clusters on)"""
    monkeypatch.setenv("VNSGPU_APPLY_PRODUCTS", "1" if request.param == "products" else "0")
    return request.param


@pytest.mark.parametrize("name", ALL)
def test_greedy_batched_builds_the_reference_tour(name, golden, eng_cache, apply_form):
    """greedy_insertion (vns_sop_solver.cpp:1091-1155) with each round as one batched evaluation + device arg-min"""
    e, sf, pb = eng_cache(name)
    g, a = golden[name], load_arrays(name)
    p = e.path()
    rounds, evals = p.greedy_insertion_batched()
    assert rounds == g["greedy"]["T"] and evals > 0
    assert p.getNumTargets() == g["greedy"]["T"] and (p.vns == a["greedy_vns"]).all()
    assert p.getPathLength() == g["greedy"]["length"] and p.getReward() == g["greedy"]["reward"]
    # both evaluation forms (endpoint vectors on / off) build the same tour
    e.set_vector_cache(0)
    try:
        q = e.path()
        q.greedy_insertion_batched()
        assert (q.vns == a["greedy_vns"]).all() and q.getPathLength() == g["greedy"]["length"]
    finally:
        e.set_vector_cache(1)


def neighbourhood(vns, T, kind):
    """every (from, to) pair insertRandom (:507-519) / exchangeRandom (:723-730) can draw, from-major; -1 rows = redrawn"""
    nvns = len(vns)
    ncol = T + 1 if kind == 1 else T
    gap = 1 if kind == 1 else 2
    fr, to = np.divmod(np.arange(nvns * ncol), ncol)
    ok = np.abs(fr - to) > gap
    inpath = fr < T
    op = np.where(inpath, OP_MOVE if kind == 1 else OP_EXCHANGE, OP_ADD if kind == 1 else OP_REPLACE)
    a = np.where(inpath, fr, vns[fr])
    return fr, to, ok, np.stack([op, a, to], 1).astype(np.int32)


def brute_force_best(pb, po, vns, T, kind, budget):
    fr, to, ok, cands = neighbourhood(vns, T, kind)
    if len(cands) == 0:
        return -1, None
    ln = np.full(len(cands), np.inf)
    ln[ok] = po.eval_batch(cands[ok])[0]
    L = po.length()
    rew = np.asarray(pb.reward, float)
    dR = np.zeros(len(cands))
    out = fr >= T
    if kind == 1:
        dR[out] = rew[vns[fr[out]]]
    else:
        dR[out] = 0.0 + (rew[vns[fr[out]]] - rew[vns[to[out]]])
    good = ok & (ln <= budget) & ((dR > 0) | ((dR == 0) & ((L - ln) > EPS)))
    if not good.any():
        return -1, None
    idx = np.flatnonzero(good)
    order = np.lexsort((idx, ln[idx], -dR[idx]))     # added reward descending, length ascending, index ascending
    k = int(idx[order[0]])
    return k, (int(fr[k]), int(to[k]), float(dR[k]), float(ln[k]))


def apply_on_vector(vns, T, kind, fr, to):
    v = list(vns)
    if kind == 2:
        v[fr], v[to] = v[to], v[fr]
        return np.array(v), T
    c = v.pop(fr)
    if fr < T:
        v.insert(to - 1 if to > fr else to, c)
        return np.array(v), T
    v.insert(to, c)
    return np.array(v), T + 1


@pytest.mark.parametrize("name", SMALL)
def test_best_move_is_the_brute_force_choice(name, golden, eng_cache, apply_form):
    from oracle import binding as ob
    e, sf, pb = eng_cache(name)
    g, a = golden[name], load_arrays(name)
    o = ob.Oracle(pb)
    vns, T = np.array(a["greedy_vns"]), g["greedy"]["T"]
    # start from a perturbed greedy tour so that both neighbourhoods hold improving moves: two targets swapped, then
    # targets dropped from the end until the tour is feasible again
    if T >= 4:
        vns[1], vns[T // 2 + 1] = vns[T // 2 + 1], vns[1]
    budget = pb.budget
    pg, po = e.path(vns, T), o.path(vns, T)
    while pg.getPathLength() > budget and T > 1:
        T -= 1
        pg.set(vns, T)
    po.set(vns, T)
    assert pg.getPathLength() <= budget
    moves = 0
    for step in range(12):
        kind = 1 + step % 2
        k, pick = brute_force_best(pb, po, vns, T, kind, budget)
        m = pg.best_move(kind)
        assert m.index == k, (step, kind, m.index, k)
        if k < 0:
            continue
        assert (m.from_, m.to) == pick[:2] and m.added_reward == pick[2] and m.length == pick[3]
        assert m.n_candidates == len(vns) * (T + 1 if kind == 1 else T)
        pg.apply_move(kind, m.from_, m.to)
        vns, T = apply_on_vector(vns, T, kind, m.from_, m.to)
        po.set(vns, T)
        assert pg.getNumTargets() == T and (pg.vns == vns).all()
        assert pg.getPathLength() == po.length() == m.length and pg.getReward() == po.reward()
        moves += 1
    assert moves > 0


@pytest.mark.parametrize("name", SMALL)
def test_best_improvement_descent(name, golden, eng_cache, apply_form):
    from oracle import binding as ob
    e, sf, pb = eng_cache(name)
    o = ob.Oracle(pb)
    p = e.path()
    p.greedy_insertion_batched()
    r0, l0 = p.getReward(), p.getPathLength()
    steps, evals = p.best_improvement()
    r1, l1 = p.getReward(), p.getPathLength()
    assert (r1, -l1) >= (r0, -l0) and l1 <= pb.budget
    assert steps == 0 or (r1, -l1) > (r0, -l0)
    assert evals > 0
    # a local optimum of both neighbourhoods, and a consistent state
    assert p.best_move(1).index == -1 and p.best_move(2).index == -1
    po = o.path(p.vns, p.getNumTargets())
    assert po.length() == l1 and po.reward() == r1
    F, B = p.layers()
    oF, oB = po.layers()[:2]
    for l in range(p.getNumTargets() + 2):
        k = pb.m[pb.start if l == 0 else (pb.goal if l == p.getNumTargets() + 1 else p.vns[l - 1])]
        assert (F[l, :k] == oF[l, :k]).all() and (B[l, :k] == oB[l, :k]).all(), l
    # the descent is deterministic
    q = e.path()
    q.greedy_insertion_batched()
    assert q.best_improvement() == (steps, evals) and (q.vns == p.vns).all()


def test_synthetic_2000_nodes_greedy_and_descent(apply_form):
    """2000 nodes / 200 clusters: the batched greedy equals the one-block greedy; the descent improves the tour"""
    from vns_sop_b200 import engine, instances as ins, synthetic as sy
    pb = ins.build_problem(sy.generate(2000, 10))
    e = engine.Engine(0).load_problem(pb)
    try:
        p, q = e.path(), e.path()
        p.greedy_insertion()
        rounds, evals = q.greedy_insertion_batched()
        assert rounds == p.getNumTargets() == q.getNumTargets() and (p.vns == q.vns).all()
        assert p.getPathLength() == q.getPathLength() and p.getReward() == q.getReward()
        r0, l0 = q.getReward(), q.getPathLength()
        steps, _ = q.best_improvement(max_steps=40)
        assert steps > 0 and (q.getReward(), -q.getPathLength()) > (r0, -l0) and q.getPathLength() <= pb.budget
    finally:
        e.close()
