mkdir -p gpurun_out
timeout 900 python bench.py --workload dop_s16 --also "" --no-e2e --cpu-sample 262144 > gpurun_out/r4g_bench_dop_s16.json 2> gpurun_out/r4g_bench_dop_s16.err; echo "exit $?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r4g_bench_dop_s16.json").read().strip().splitlines()[-1])
print("value %.4g" % d["value"]); t = d.get("time_to_reward", {})
print({k: t.get(k) for k in ("gpu_wall_s", "cpu_reference_wall_s", "gpu_time_to_final_reward_s", "cpu_time_to_final_reward_s", "same_result", "reward")})
print(json.dumps(t.get("multi_start"))[:400]); print(json.dumps(t.get("race_to_reference_reward"))[:500])
print(json.dumps(d.get("full_neighbourhood"))[:300])
PY
