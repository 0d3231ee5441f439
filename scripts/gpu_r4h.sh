# r4h: third chain form with the table in global memory: parity of the new modes, dop_s16 (config 3) solve legs, krob200 with the
# table forced to global memory for comparison
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "smem-tab" > gpurun_out/r4h_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r4h_tests.log
tail -5 gpurun_out/r4h_tests.log
timeout 900 python bench.py --workload dop_s16 --also "" --no-e2e --cpu-sample 262144 > gpurun_out/r4h_bench_dop_s16.json 2> gpurun_out/r4h_bench_dop_s16.err; echo "exit $?"
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r4h_bench_dop_s16.json").read().strip().splitlines()[-1])
t = d.get("time_to_reward", {})
print("dop_s16", {k: t.get(k) for k in ("gpu_wall_s", "cpu_reference_wall_s", "gpu_time_to_final_reward_s", "cpu_time_to_final_reward_s", "same_result", "reward")})
print(json.dumps(t.get("multi_start"))[:300])
PY
for v in 2 3; do VNSGPU_CHAIN3_U16=$v timeout 300 python scripts/time_solve.py krob200 0 3 >> gpurun_out/r4h_solve_krob200_global.jsonl 2>> gpurun_out/r4h_solve.err; done
cat gpurun_out/r4h_solve_krob200_global.jsonl | cut -c1-260
