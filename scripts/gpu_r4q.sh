# r4q: default bench line with config 3's solve legs
mkdir -p gpurun_out
( time timeout 1500 python bench.py > gpurun_out/r4q_bench.json 2> gpurun_out/r4q_bench.err ) 2> gpurun_out/r4q_bench.time; echo "bench exit $?"; tail -3 gpurun_out/r4q_bench.err
cat gpurun_out/r4q_bench.time
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r4q_bench.json").read().strip().splitlines()[-1])
print("value %.4g e2e %.4g" % (d["value"], d["e2e"]["value"]))
for k in ("time_to_reward", "time_to_reward_dop_s16"):
    t = d[k]; print(k, {q: t.get(q) for q in ("gpu_wall_s", "cpu_reference_wall_s", "same_result", "reward")}, "multi", t["multi_start"]["gpu_runs_per_s"], t["multi_start"].get("cpu_runs_per_s"))
print("vm1084", d["time_to_reward_vm1084"]["gpu_wall_s"])
PY
