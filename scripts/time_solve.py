"""Seeded full VNS solve (stock sop.cfg stopping rule) on one instance: wall time, trajectory against the golden record."""
import json, sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from conftest import load_problem
from vns_sop_b200 import engine
name = sys.argv[1] if len(sys.argv) > 1 else "krob200"
blocks = int(sys.argv[2]) if len(sys.argv) > 2 else 0
form = int(sys.argv[3]) if len(sys.argv) > 3 else -1
sf, pb = load_problem(name)
e = engine.Engine(0).load_problem(pb)
e.set_chain_blocks(blocks)
e.set_chain_form(form)
g = json.load(open("tests/golden/golden.json"))[name].get("solve_seed1")
p = e.path()
e.sync()
t0 = time.perf_counter()
res, impr = p.vns_run(1)
dt = time.perf_counter() - t0
print(json.dumps({"instance": name, "chain_blocks": blocks, "chain_form": form, "wall_s": dt, "device_ms": res.time_ms, "reward": res.reward, "length": res.length,
                  "num_iters": res.num_iters, "last_impr": res.num_iters_last_impr, "proposals": res.proposals, "evals": res.evals,
                  "accepts": res.accepts, "time_to_final_reward_s": res.time_last_impr_ms / 1e3,
                  "same_as_reference": None if not g else bool(res.reward == g["REWARDS"] and res.length == g["LENGTH"] and res.num_iters == g["NUM_ITERS"]
                                                                and [[r, l] for r, l, _ in impr] == g["IMPROVEMENTS"])}), flush=True)
