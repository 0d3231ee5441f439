mkdir -p gpurun_out
( time timeout 1500 python bench.py > gpurun_out/r4v_bench.json 2> gpurun_out/r4v_bench.err ) 2> gpurun_out/r4v_bench.time; echo "bench exit $?"; tail -2 gpurun_out/r4v_bench.err; grep real gpurun_out/r4v_bench.time
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r4v_bench.json").read().strip().splitlines()[-1])
def show(name, fn):
    b = fn["best_improvement"]; g = fn["greedy_batched"]
    print(name, "greedy %.3f s reward %g | descent %.3f s %d moves reward %g length %g | exact chain" % (g["seconds"], g["reward"], b["seconds"], b["moves"], b["reward"], b["length"]), {k: b["exact_chain_same_time"][k] for k in ("seconds", "reward", "length", "vns_steps", "accepts")})
show("krob200", d["full_neighbourhood"])
for w in ("vm1084", "synth20k"): show(w, d["workloads"][w]["full_neighbourhood"])
print("value %.4g e2e %.4g ttr %.3f" % (d["value"], d["e2e"]["value"], d["time_to_reward"]["gpu_wall_s"]))
PY
