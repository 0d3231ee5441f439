# r4r: third chain form at 128 threads per chain: parity of every chain mode, seeded solves, dense multi-start, default bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_cli.py -m gpu -x -q -k "sequence or edge_cases or vns_run or cli" > gpurun_out/r4r_chain_tests.log 2>&1; echo "pytest exit $?" >> gpurun_out/r4r_chain_tests.log
tail -4 gpurun_out/r4r_chain_tests.log
for inst in krob200 dop_s12 opn_s08 berlin52; do timeout 300 python scripts/time_solve.py $inst 0 -1 >> gpurun_out/r4r_solve_small.jsonl 2>/dev/null; done; cut -c1-150 gpurun_out/r4r_solve_small.jsonl
( time timeout 1500 python bench.py > gpurun_out/r4r_bench.json 2> gpurun_out/r4r_bench.err ) 2> gpurun_out/r4r_bench.time; echo "bench exit $?"; tail -2 gpurun_out/r4r_bench.err; grep real gpurun_out/r4r_bench.time
python - <<'PY'
import json
d = json.loads(open("gpurun_out/r4r_bench.json").read().strip().splitlines()[-1])
for k in ("time_to_reward", "time_to_reward_dop_s16"):
    t = d[k]; m = t["multi_start"]
    print(k, {q: t.get(q) for q in ("gpu_wall_s", "cpu_reference_wall_s", "same_result")}, "multi %.1f runs/s" % m["gpu_runs_per_s"], "dense", json.dumps(m.get("dense"))[:260], "race %.3f" % t["race_to_reference_reward"]["gpu_wall_s"])
PY
