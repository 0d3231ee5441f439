# r4j: both bench arms at the round's last kernel commit
mkdir -p gpurun_out
( time timeout 1500 python bench.py > gpurun_out/r4j_bench.json 2> gpurun_out/r4j_bench.err ) 2> gpurun_out/r4j_bench.time; echo "bench exit $?"
timeout 600 python bench.py --impl reference > gpurun_out/r4j_bench_ref.json 2> gpurun_out/r4j_bench_ref.err; echo "ref exit $?"
cat gpurun_out/r4j_bench.time
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r4j_bench.json").read().strip().splitlines()[-1])
r = json.loads(open("gpurun_out/r4j_bench_ref.json").read().strip().splitlines()[-1])
print("value %.4g e2e %.4g ref %.4g (%s cores) -> e2e ratio %.0f" % (d["value"], d["e2e"]["value"], r["value"], r["cpu_baseline"]["cores"], d["e2e"]["value"] / r["value"]))
t = d["time_to_reward"]; print("ttr gpu %.3f cpu %.3f same %s" % (t["gpu_wall_s"], t["cpu_reference_wall_s"], t["same_result"]))
print("roofline", {k: d["roofline"][k] for k in ("kernel", "bound", "frac", "traffic")})
print("roofline_hbm", {k: d["roofline_hbm"].get(k) for k in ("kernel", "bound", "frac", "traffic", "algorithmic_bytes_per_launch")})
fn = d["workloads"]["synth20k"]["full_neighbourhood"]; print("synth20k greedy %.2f s, descent %.2f s" % (fn["greedy_batched"]["seconds"], fn["best_improvement"]["seconds"]))
print("clocks", d["clocks"])
PY
