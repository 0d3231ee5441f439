"""Time vnsgpu_local_search on the synthetic scaled instances (config 5) per chain configuration."""
import sys, time
sys.path.insert(0, "."); sys.path.insert(0, "tests")
from vns_sop_b200 import engine, instances as ins, synthetic as sy
nodes = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
modes = sys.argv[2].split(",") if len(sys.argv) > 2 else ["1t", "8t", "0t"]
nprop = int(sys.argv[3]) if len(sys.argv) > 3 else 200000
sf = sy.generate(nodes, 10) if nodes != 20000 else sy.generate()
pb = ins.build_problem(sf)
e = engine.Engine(0).load_problem(pb)
vns, T = sy.centroid_tour_state(pb)
for mode in modes:     # <blocks>[t]: cooperating blocks (0 = default), t = product tree on
    e.set_chain_blocks(int(mode.rstrip("t")))
    e.set_chain_tree(1 if mode.endswith("t") else 0)
    p = e.path(vns, T); p.fitDOPtoBudget(); rng = engine.seeded_rng(7)
    e.chain_cycles(True)
    for k in (1, 2):
        p.shake(k, rng); e.sync()
        t0 = time.perf_counter(); st = p._ls(k, nprop, rng, None, 1); dt = time.perf_counter() - t0
        print("synth%d %-4s kind %d: T=%d %8d proposals %6d accepts %6d rounds  %.3f s  %.3f us/proposal  %.1f us/round"
              % (nodes, mode, k, p.getNumTargets(), st.proposals, st.accepts, st.rounds, dt, 1e6 * dt / st.proposals, 1e6 * dt / max(st.rounds, 1)), flush=True)
    print("   cycles", e.chain_cycles(), flush=True)
