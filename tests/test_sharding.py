"""N > 1 host path on CPU: two gloo ranks shard one candidate batch, evaluate their shares with the
CPU oracle standing in for the per-rank engine, and reduce the chosen move; the result must equal
the single-rank answer (and the shards must tile the batch)."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

from conftest import load_arrays, load_problem, random_candidates

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_shard_range_tiles_the_batch():
    from vns_sop_b200.sharding import shard_range
    for n in (0, 1, 7, 64, 1000003):
        for world in (1, 2, 3, 8):
            r = [shard_range(n, k, world) for k in range(world)]
            assert r[0][0] == 0 and r[-1][1] == n
            assert all(r[k][1] == r[k + 1][0] for k in range(world - 1))
            sizes = [hi - lo for lo, hi in r]
            assert max(sizes) - min(sizes) <= 1


def _oracle_eval(name, golden_T, golden_vns):
    from oracle import binding as ob
    sf, pb = load_problem(name)
    o = ob.Oracle(pb)
    p = o.path(golden_vns, golden_T)
    L, budget = p.length(), pb.budget

    def fn(c):
        ln, lb = p.eval_batch(c)
        dR = np.zeros(len(c))
        add, rep = c[:, 0] == 0, c[:, 0] == 1
        dR[add] = pb.reward[c[add, 1]]
        dR[rep] = pb.reward[c[rep, 1]] - pb.reward[golden_vns[c[rep, 2]]]
        acc = (ln <= budget) & ((dR > 0) | ((L - ln) > 1e-5)) & ~((lb > budget) & (c[:, 0] != 2))
        idx = np.nonzero(acc)[0]
        return ln, lb, acc.astype(np.uint8) * 2, (int(idx[0]) if len(idx) else -1)
    return fn


def _worker(rank, world, port, name, T, vns, cands, out):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import torch.distributed as dist
    from vns_sop_b200 import sharding
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    fn = _oracle_eval(name, T, vns)
    lo, hi, ln, lb, fl, first = sharding.sharded_first_accept(fn, cands, rank, world)
    # the device-side variant used inside bench.py's timed loop
    import torch
    local = fn(cands[lo:hi])[3]
    key = sharding.reduce_first_accept_(torch.tensor([local], dtype=torch.int64), lo)
    # the sharded entry point (vnsgpu_eval_batch_dev_shard) writes index_base + local index or NO_KEY itself
    from vns_sop_b200 import engine
    key2 = sharding.reduce_key_(torch.tensor([engine.NO_KEY if local < 0 else lo + local], dtype=torch.int64))
    out.put((rank, lo, hi, first, int(key.item()), float(ln.sum()), int(key2.item())))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("case", ["accepting", "none-accepted"])
def test_two_rank_first_accept_equals_single_rank(case, golden):
    from vns_sop_b200 import sharding
    name = "berlin52"
    g, a = golden[name], load_arrays(name)
    T, vns = g["greedy"]["T"], a["greedy_vns"]
    rs = np.random.RandomState(4)
    cands = random_candidates(rs, vns, T, 4001)
    fn = _oracle_eval(name, T, vns)
    ln, lb, fl, first1 = fn(cands)
    if case == "none-accepted":
        cands = cands[(fl & 2) == 0]            # keep only rejected candidates
        ln, lb, fl, first1 = fn(cands)
        assert first1 == -1
    else:
        # put the first accepted candidate into the SECOND rank's shard
        order = np.argsort((fl & 2) != 0, kind="stable")
        cands = cands[order]
        ln, lb, fl, first1 = fn(cands)
        assert first1 >= sharding.shard_range(len(cands), 1, 2)[0]
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, name, T, vns, cands, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(out.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][1] == 0 and res[0][2] == res[1][1] and res[1][2] == len(cands)
    assert res[0][3] == res[1][3] == first1
    want_key = sharding.NONE if first1 < 0 else first1
    assert res[0][4] == res[1][4] == want_key
    assert abs(res[0][5] + res[1][5] - float(ln.sum())) < 1e-6
    from vns_sop_b200 import engine
    assert res[0][6] == res[1][6] == (engine.NO_KEY if first1 < 0 else first1)
