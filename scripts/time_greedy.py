"""Batched greedy_insertion + best-improvement descent on one bench workload: wall time per round / move."""
import sys, time
sys.path.insert(0, ".")
import bench
from vns_sop_b200 import engine
name = sys.argv[1] if len(sys.argv) > 1 else "synth20k"
sf, pb = bench.load_workload(name)
e = engine.Engine(0).load_problem(pb)
p = e.path(); e.sync()
t0 = time.perf_counter(); rounds, evals = p.greedy_insertion_batched(); e.sync(); t1 = time.perf_counter()
print("%s greedy_batched: %d rounds, %d candidates, %.3f s, %.2f ms / round, %.1f M evals/s; T=%d reward=%g length=%g"
      % (name, rounds, evals, t1 - t0, 1e3 * (t1 - t0) / max(rounds, 1), evals / (t1 - t0) / 1e6, p.getNumTargets(), p.getReward(), p.getPathLength()), flush=True)
t1 = time.perf_counter(); steps, ev2 = p.best_improvement(max_steps=100); e.sync(); t2 = time.perf_counter()
print("%s best_improvement: %d moves, %d candidates, %.3f s, %.2f ms / move, %.1f M evals/s; T=%d reward=%g length=%g"
      % (name, steps, ev2, t2 - t1, 1e3 * (t2 - t1) / max(steps, 1), ev2 / (t2 - t1) / 1e6, p.getNumTargets(), p.getReward(), p.getPathLength()), flush=True)
