"""sop_vns_b200 - the reference's `sop_vns --gui=none --problem=X.sop` front end over the GPU engine
(vns-sop_b200/host/).  CPU tests: the C++ loader against the Python restatement of the reference loader
(which is pinned to the compiled reference by tests/golden/make_golden.py), option handling, and the loud
failure without a GPU.  GPU tests: a seeded solve writes the same results.log record and path.txt as the
reference binary (every column but the wall-clock ones)."""
import json
import os
import re
import subprocess

import numpy as np
import pytest

from conftest import ALL, GOLDEN_DIR, ROOT, SMALL

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "sop_vns_ref")
TIMED = ("CTIME", "CTIME_LAST_IMPR")


@pytest.fixture(scope="session")
def cli():
    from vns_sop_b200 import build as b
    b.build()
    return b.build_cli()


def write_fixture(name, tmp_path):
    from vns_sop_b200 import instances as ins
    sf = ins.load_npz(os.path.join(GOLDEN_DIR, name + ".npz"))
    f = str(tmp_path / (name + ".sop"))
    ins.write_sop(sf, f)
    return sf, f


def parse_record(line):
    rec = {}
    for item in line.strip().split(";"):
        if item:
            k, _, v = item.partition(":")
            rec[k] = v
    return rec


def untimed(rec):
    r = {k: v for k, v in rec.items() if k not in TIMED and k != "PROBLEM_FILE"}
    r["SOLUTION_TIME_REWARD_LENGTH_IMPROVEMENTS"] = [x.split("|")[1:] for x in rec["SOLUTION_TIME_REWARD_LENGTH_IMPROVEMENTS"].split(",") if x]
    return r


@pytest.mark.parametrize("name", ALL)
def test_cpp_loader_equals_python_loader(cli, name, tmp_path):
    from vns_sop_b200 import instances as ins
    sf, f = write_fixture(name, tmp_path)
    pb = ins.build_problem(ins.parse_sop(f))
    out = subprocess.run([cli, "--problem=" + f, "--dump-problem=1"], capture_output=True, text=True, check=True).stdout
    d = json.loads(out)
    assert d["C"] == pb.C and d["N"] == pb.N and d["start"] == pb.start and d["goal"] == pb.goal
    assert d["budget"] == pb.budget and d["rule"] == pb.rule and d["name"] == pb.name
    assert d["m"] == pb.m.tolist() and d["reward"] == pb.reward.tolist()
    assert d["node_id"] == pb.node_id.tolist() and d["orig_cluster_id"] == pb.orig_cluster_id.tolist()
    if pb.xy is not None:
        assert d["xy"] == pb.xy.ravel().tolist() and d["W_size"] == 0
    else:
        assert d["W_size"] == pb.N * pb.N and d["W_sum"] == float(pb.W.sum())
    assert d["max_reward"] == float(pb.reward.sum())


def test_cpp_loader_on_the_reference_dataset_files(cli):
    """Every .sop file of the reference's sop/ set and a sample of the OPN / DOP sets (start != 0,
    start != goal and EXPLICIT cases included)."""
    from vns_sop_b200 import instances as ins
    base = "/root/reference/datasets"
    if not os.path.isdir(base):
        pytest.skip("reference datasets not present on this machine")
    files = []
    for sub, every in (("sop", 4), ("opn_sop_dataset", 12), ("dop_sop_dataset", 12)):
        d = os.path.join(base, sub)
        if os.path.isdir(d):
            fs = sorted(x for x in os.listdir(d) if x.endswith(".sop"))
            files += [os.path.join(d, x) for x in fs[::every] if os.path.getsize(os.path.join(d, x)) < 3e6]
    assert len(files) >= 10
    for f in files:
        pb = ins.build_problem(ins.parse_sop(f))
        d = json.loads(subprocess.run([cli, "--problem=" + f, "--dump-problem=1"], capture_output=True, text=True, check=True).stdout)
        assert d["m"] == pb.m.tolist() and d["node_id"] == pb.node_id.tolist(), f
        assert d["orig_cluster_id"] == pb.orig_cluster_id.tolist() and d["reward"] == pb.reward.tolist(), f
        assert d["budget"] == pb.budget and d["rule"] == pb.rule and d["goal"] == pb.goal, f


def test_cli_options(cli, tmp_path):
    r = subprocess.run([cli, "--bogus=1"], capture_output=True, text=True)
    assert r.returncode == 2 and "unrecognised option" in r.stderr
    r = subprocess.run([cli], capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode == 2 and "Problem file must be specified" in r.stderr
    r = subprocess.run([cli, "--problem=x.sop", "--gui=cairo"], capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode == 2 and "headless" in r.stderr
    r = subprocess.run([cli, "--problem=/nonexistent.sop"], capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode == 1 and "Cannot open problem file" in r.stderr
    # sop.cfg in the working directory is read; the command line wins (sop.cpp:224-233); GUI keys are accepted
    _, f = write_fixture("berlin52", tmp_path)
    (tmp_path / "sop.cfg").write_text("#comment\ngui=none\nproblem=/nonexistent.sop\ndraw-path=1\nnum-iterations=7\n")
    r = subprocess.run([cli, "--problem=" + f, "--dump-problem=1"], capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode == 0 and json.loads(r.stdout)["C"] == 13
    (tmp_path / "bad.cfg").write_text("no-such-key=1\n")
    r = subprocess.run([cli, "--problem=" + f, "--config-file=bad.cfg"], capture_output=True, text=True, cwd=str(tmp_path))
    assert r.returncode == 1 and "unrecognised option" in r.stderr


def test_cli_without_gpu_fails_loudly(cli, tmp_path):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    _, f = write_fixture("berlin52", tmp_path)
    r = subprocess.run([cli, "--problem=" + f, "--seed=1", "--output=" + str(tmp_path / "o")], capture_output=True, text=True)
    assert r.returncode == 1 and "vnsgpu" in r.stderr
    assert not (tmp_path / "o" / "results.log").exists()


def _solve(cli, f, out, *extra):
    subprocess.run([cli, "--problem=" + f, "--output=" + out] + list(extra), capture_output=True, text=True, check=True)
    with open(os.path.join(out, "results.log")) as fh:
        return [parse_record(l) for l in fh if l.strip()]


@pytest.mark.gpu
@pytest.mark.parametrize("name,loop", [("berlin52", 1), ("berlin52", 0), ("opn_s08", 1), ("dop_s12", 1), ("dop_s12", 0), ("krob200", 1)])
def test_seeded_solve_writes_the_reference_record(cli, golden, name, loop, tmp_path):
    sf, f = write_fixture(name, tmp_path)
    out = str(tmp_path / "gpu")
    rec = _solve(cli, f, out, "--seed=1", "--device-loop=%d" % loop)[0]
    s = golden[name]["solve_seed1"]
    assert int(rec["NUM_ITERS"]) == s["NUM_ITERS"] and int(rec["NUM_ITERS_LAST_IMPR"]) == s["NUM_ITERS_LAST_IMPR"]
    assert float(rec["REWARDS"]) == s["REWARDS"] and float(rec["LENGTH"]) == s["LENGTH"] and float(rec["BUDGET"]) == s["BUDGET"]
    assert float(rec["MAX_ACHIEVABLE_REWARDS"]) == s["MAX_ACHIEVABLE_REWARDS"]
    assert [int(x) for x in rec["RESULT_TARGET_IDS"].split(",")] == s["RESULT_TARGET_IDS"]
    assert [int(x) for x in rec["RESULT_CLUSTER_IDS"].split(",")] == s["RESULT_CLUSTER_IDS"]
    impr = [[float(y) for y in x.split("|")[1:]] for x in rec["SOLUTION_TIME_REWARD_LENGTH_IMPROVEMENTS"].split(",")]
    assert impr == s["IMPROVEMENTS"]
    assert rec["METHOD"] == "vns_sop" and rec["NAME"] == sf.name
    for fn in ("path.txt", "gop.txt", "targets.txt", "settings.txt", "info.txt"):
        assert os.path.getsize(os.path.join(out, fn)) > 0
    if os.path.exists(REF_BIN):   # byte-level comparison with the compiled reference's own files
        ref_out = str(tmp_path / "ref")
        subprocess.run([REF_BIN, "--problem=" + f, "--seed=1", "--output=" + ref_out], capture_output=True, check=True)
        with open(os.path.join(ref_out, "results.log")) as fh:
            ref = parse_record(fh.readline())
        assert untimed(rec) == untimed(ref)
        assert list(rec.keys()) == list(ref.keys())
        for fn in ("path.txt", "gop.txt", "targets.txt"):
            assert open(os.path.join(out, fn)).read() == open(os.path.join(ref_out, fn)).read(), fn


@pytest.mark.gpu
def test_batch_mode_one_launch_per_batch(cli, golden, tmp_path):
    """--batch=N: N records, iteration directories 0..N-1, run i seeded seed + i (run 0 = the golden seed-1 solve)."""
    _, f = write_fixture("berlin52", tmp_path)
    out = str(tmp_path / "b")
    recs = _solve(cli, f, out, "--seed=1", "--batch=5")
    assert len(recs) == 5
    s = golden["berlin52"]["solve_seed1"]
    assert float(recs[0]["REWARDS"]) == s["REWARDS"] and float(recs[0]["LENGTH"]) == s["LENGTH"]
    assert [int(x) for x in recs[0]["RESULT_TARGET_IDS"].split(",")] == s["RESULT_TARGET_IDS"]
    for i in range(5):
        assert os.path.getsize(os.path.join(out, str(i), "path.txt")) > 0
    singles = [_solve(cli, f, str(tmp_path / ("s%d" % i)), "--seed=%d" % (1 + i))[0] for i in (1, 3)]
    for rec, i in zip(singles, (1, 3)):
        assert rec["RESULT_TARGET_IDS"] == recs[i]["RESULT_TARGET_IDS"] and rec["LENGTH"] == recs[i]["LENGTH"]
