"""Time vnsgpu_local_search (insertRandom / exchangeRandom) per chain implementation on one instance."""
import sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from conftest import load_problem
from vns_sop_b200 import engine
name = sys.argv[1] if len(sys.argv) > 1 else "krob200"
modes = sys.argv[2].split(",") if len(sys.argv) > 2 else ["c2-1", "c2-4", "v1"]
sf, pb = load_problem(name)
e = engine.Engine(0).load_problem(pb)
for mode in modes:
    e.set_chain_form(3 if mode == "c3" else (1 if mode == "v1" else 0))
    e.set_chain_blocks(int(mode.split("-")[1].rstrip("t")) if mode.startswith("c2-") else 1)
    e.set_chain_tree(1 if mode.endswith("t") and mode != "v1" else 0)
    p = e.path(); p.greedy_insertion(); rng = engine.seeded_rng(7)
    n = len(p.vns) ** 2
    tot = {1: [0, 0, 0, 0.0], 2: [0, 0, 0, 0.0]}
    e.chain_cycles(True)
    for it in range(20 if name != "vm1084" else 3):
        for k in (1, 2):
            p.shake(k, rng); e.sync()
            t0 = time.perf_counter(); st = p._ls(k, n, rng, None, 1); dt = time.perf_counter() - t0
            t = tot[k]; t[0] += st.proposals; t[1] += st.accepts; t[2] += st.rounds; t[3] += dt
    for k in (1, 2):
        t = tot[k]
        print("%s %-6s kind %d: %9d proposals %6d accepts %7d rounds  %.3f s  %.3f us/proposal  %.2f us/round"
              % (name, mode, k, t[0], t[1], t[2], t[3], 1e6 * t[3] / max(t[0], 1), 1e6 * t[3] / max(t[2], 1)), flush=True)
    print("   cycles", e.chain_cycles())
    print("   final T=%d length=%g reward=%g" % (p.getNumTargets(), p.getPathLength(), p.getReward()), flush=True)
