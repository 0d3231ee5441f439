"""Two-GPU paths (skipped on a one-GPU box; run with `gpurun --gpus 2`): the NCCL exchange steps behind the C ABI
(vnsgpu_comm_init from a unique id, vnsgpu_path_broadcast, vnsgpu_eval_batch_dev_shard + vnsgpu_reduce_first_key) and the
command line's replica mode (--gpus=N --batch=M)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from conftest import load_problem, random_candidates

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


needs2 = pytest.mark.skipif(_gpus() < 2, reason="needs two GPUs")

_WORKER = r'''
import os, sys, time
import numpy as np
sys.path.insert(0, %(root)r); sys.path.insert(0, os.path.join(%(root)r, "tests"))
import torch
from conftest import load_problem, random_candidates
from vns_sop_b200 import engine
rank, world, tmp = int(sys.argv[1]), int(sys.argv[2]), sys.argv[3]
sf, pb = load_problem("krob200")
torch.cuda.set_device(rank)
e = engine.Engine(rank).load_problem(pb)
idf = os.path.join(tmp, "id.bin")
if rank == 0:
    with open(idf + ".tmp", "wb") as fh:
        fh.write(e.comm_unique_id())
    os.rename(idf + ".tmp", idf)
while not os.path.exists(idf):
    time.sleep(0.05)
e.comm_init(open(idf, "rb").read(), rank, world)
p = e.path()
if rank == 0:                      # only the root holds the incumbent
    p.greedy_insertion()
    rng = engine.seeded_rng(3)
    p.shake(1, rng)
    p._ls(1, 400, rng, None, 1)
e._ck(e.lib.vnsgpu_path_broadcast(e.h, p.h, 0, None))
e.sync()
rs = np.random.RandomState(8)
T, vns = p.getNumTargets(), np.array(p.vns)
cands = p.make_cands(random_candidates(rs, vns, T, 40000))     # same list on both ranks (same seed, same incumbent)
lo, hi = rank * len(cands) // world, (rank + 1) * len(cands) // world
dev = torch.device("cuda", rank)
d_c = torch.from_numpy(cands[lo:hi].view(np.uint8).copy()).to(dev)
d_key = torch.zeros(1, dtype=torch.int64, device=dev)
torch.cuda.synchronize()
import ctypes as C
e._ck(e.lib.vnsgpu_eval_batch_dev_shard(e.h, p.h, hi - lo, C.c_void_p(d_c.data_ptr()), pb.budget, 1, None, None, None,
                                        C.c_void_p(d_key.data_ptr()), lo, None))
e.reduce_first_key(d_key.data_ptr())
e.sync()
ln, lb, fl, first = p.eval_batch(cands)                        # the whole batch on this rank alone
# full-neighbourhood modes: every neighbourhood split over the two ranks = the single-GPU search
q1 = e.path(); q1.greedy_insertion_batched(); s1 = q1.best_improvement()
e.set_shard_neighbourhoods(1)
q2 = e.path(); r2 = q2.greedy_insertion_batched(); s2 = q2.best_improvement()
m = q2.best_move(1)
e.set_shard_neighbourhoods(0)
np.savez(os.path.join(tmp, "r%%d.npz" %% rank), T=T, vns=vns, length=p.getPathLength(), key=int(d_key.item()), first=first,
         nb_same=bool((q1.vns == q2.vns).all() and q1.getPathLength() == q2.getPathLength() and q1.getReward() == q2.getReward()),
         nb_vns=np.array(q2.vns), nb_steps=np.array([s1[0], s2[0]]), nb_evals=np.array([s1[1], s2[1]]), nb_last=m.index)
e.comm_destroy()
e.close()
'''


@needs2
@pytest.mark.gpu
def test_comm_broadcast_and_first_key_two_ranks(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(_WORKER % {"root": ROOT})
    procs = [subprocess.Popen([sys.executable, str(script), str(r), "2", str(tmp_path)]) for r in range(2)]
    assert [q.wait(timeout=600) for q in procs] == [0, 0]
    a, b = np.load(tmp_path / "r0.npz"), np.load(tmp_path / "r1.npz")
    assert int(a["T"]) == int(b["T"]) > 0 and (a["vns"] == b["vns"]).all() and float(a["length"]) == float(b["length"])   # the incumbent arrived
    assert int(a["key"]) == int(b["key"]) == int(a["first"]) == int(b["first"])      # the reduced key = the single-GPU answer
    assert int(a["first"]) >= 0
    # sharded full-neighbourhood descent: same tour as unsharded on both ranks, the same number of moves, and the two
    # shards' candidate counts add up to the whole neighbourhoods
    assert bool(a["nb_same"]) and bool(b["nb_same"]) and (a["nb_vns"] == b["nb_vns"]).all()
    assert a["nb_steps"][0] == a["nb_steps"][1] == b["nb_steps"][1]
    assert int(a["nb_evals"][1]) + int(b["nb_evals"][1]) == int(a["nb_evals"][0])
    assert int(a["nb_last"]) == int(b["nb_last"]) == -1


@needs2
@pytest.mark.gpu
def test_cli_batch_over_two_gpus_equals_one(tmp_path):
    from test_cli import _solve, write_fixture
    from vns_sop_b200 import build as b
    cli = b.build_cli()
    _, f = write_fixture("berlin52", tmp_path)
    one = _solve(cli, f, str(tmp_path / "g1"), "--seed=1", "--batch=6")
    two = _solve(cli, f, str(tmp_path / "g2"), "--seed=1", "--batch=6", "--gpus=2")
    assert len(one) == len(two) == 6
    for x, y in zip(one, two):
        for k in ("REWARDS", "LENGTH", "NUM_ITERS", "RESULT_TARGET_IDS", "RESULT_CLUSTER_IDS"):
            assert x[k] == y[k], k
